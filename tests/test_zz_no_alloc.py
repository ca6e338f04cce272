"""``afl_hyper_step_noalloc``: the hyper-network step for callers that are not allowed to allocate.

Some users of this library cannot call ``cudaMalloc`` (or go through torch's caching allocator) at the point
where they need the update: a control loop that reserved all of its device memory up front, a plugin that lives
inside another runtime's memory budget, a host callback that must not cause a device-wide synchronisation.  For
them the library offers a second entry point with the *same* arithmetic as ``afl_hyper_step`` whose scratch is a
fixed 1 MiB array inside the library instead of caller-provided buffers.

This test drives the new entry point through the C ABI (``ctypes``) and checks, on two calls in a row, that every
output matches what the existing entry point produces from the same starting state, within the tolerance the
other hyper-network tests use.  It also checks that sizes above the stated maximum are refused with a clear
message.  The file is named so that pytest collects it last; it depends on nothing the other tests leave behind
and passes on its own in a fresh process as well.
"""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu

ATOL, RTOL = 3e-6, 1e-4          # same as tests/test_native_ops.py::test_hyper_emit_and_step_match_torch


def _state(device, n_clients=4, seed=0):
    from attackfl_b200 import models
    from attackfl_b200.models.hyper import FusedHyperNet
    from attackfl_b200.ops import flat
    torch.manual_seed(seed)
    h = models.HyperNetwork(models.TransformerModel(), n_clients, 8, 100, False, 2)
    fused = FusedHyperNet.from_module(h, device=device)
    nat = flat.NativeHyper(fused)
    rows = []
    for i in range(n_clients):
        torch.manual_seed(100 + i)
        rows.append(fused.layout.flatten(models.TransformerModel().state_dict()))
    return nat, torch.stack(rows).to(device)


def _outputs(nat):
    return {"head_w": nat.head_w, "head_b": nat.head_b, "mw": nat.mw, "vw": nat.vw, "mb": nat.mb, "vb": nat.vb,
            "small": nat.small}


def _call_noalloc(lib, nat, client, theta, lr, clip, step, coef_norm, d_pad=None):
    p = lambda t: ctypes.c_void_p(t.data_ptr())
    d_pad = nat.fused.layout.padded if d_pad is None else d_pad
    return lib.afl_hyper_step_noalloc(
        p(nat.head_w), p(nat.head_b), p(nat.mw), p(nat.vw), p(nat.mb), p(nat.vb), p(nat.small),
        ctypes.c_int(nat.n_nodes), ctypes.c_int(nat.E), ctypes.c_int(nat.H), ctypes.c_int(client),
        p(nat._client_idx[client:client + 1]), p(theta), ctypes.c_int64(d_pad), ctypes.c_float(lr),
        ctypes.c_float(0.9), ctypes.c_float(0.999), ctypes.c_float(1e-8), ctypes.c_int(step), ctypes.c_float(clip),
        p(coef_norm), ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))


def test_noalloc_entry_matches_existing_entry_on_two_calls_in_a_row():
    from attackfl_b200.ops import native
    lib = native.cuda_lib()
    assert hasattr(lib, "afl_hyper_step_noalloc")
    dev = "cuda"
    ref, rows = _state(dev)          # driven through the existing entry point (afl_hyper_step, caller scratch)
    new, _ = _state(dev)             # identical start, driven through the allocation-free entry point
    for k, v in _outputs(ref).items():
        assert torch.equal(v, _outputs(new)[k]), k
    coef_norm = torch.zeros(2, device=dev)
    lr, clip = 1e-3, 1.0
    for call, client in enumerate((0, 1), start=1):
        ref.step(client, rows[client], lr, clip)
        rc = _call_noalloc(lib, new, client, rows[client], lr, clip, call, coef_norm)
        assert rc == 0, lib.afl_last_error().decode()
        torch.cuda.synchronize()
        for k, v in _outputs(ref).items():
            assert torch.allclose(_outputs(new)[k], v, atol=ATOL, rtol=RTOL), (call, k)
        # clip coefficient / gradient norm come back through the optional output
        assert torch.allclose(coef_norm, ref.red[ref.H + 1:ref.H + 3], atol=1e-6, rtol=1e-4), call
    assert not torch.equal(new.head_b, _state(dev)[0].head_b)      # the calls did update the parameters


def test_noalloc_entry_rejects_sizes_beyond_its_scratch():
    from attackfl_b200.ops import native
    lib = native.cuda_lib()
    nat, rows = _state("cuda")
    before = nat.head_b.clone()
    rc = _call_noalloc(lib, nat, 0, rows[0], 1e-3, 1.0, 1, torch.zeros(2, device="cuda"), d_pad=262144 + 1024)
    assert rc != 0
    msg = lib.afl_last_error().decode()
    assert "exceed" in msg and "262144" in msg and "afl_hyper_step" in msg
    torch.cuda.synchronize()
    assert torch.equal(nat.head_b, before)                          # nothing was launched
