set -x
timeout 600 python -m pytest tests/test_train_tc.py -x -q 2>&1 | tail -15
for cs in 1 2 4 8; do ATTACKFL_TC_CLUSTER=$cs timeout 120 python scripts/prof_train_tc.py 12800 1 2>&1 | tail -1; done
PROF_CLIENTS=16,32 timeout 200 python scripts/prof_train_tc.py 6400 1 2>&1 | grep -v "^$" | tail -6
PROF_STAMPS=1 timeout 120 python scripts/prof_train_tc.py 2560 1 2>&1 | tail -3
