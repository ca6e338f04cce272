from .core import (GAMMA0, TAU, SearchResult, craft, gamma_schedule, lie_attack, mean_std,  # noqa: F401
                   min_max_attack, min_sum_attack, opt_fang_attack, random_attack,
                   reference_distance)
