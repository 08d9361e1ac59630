"""ngp_nerf_network_inference_noalloc (include/ngp_b200.h; csrc/noalloc.cu): the allocation-free variant of NerfNetwork inference
(no counterpart in the reference).  Called through the C ABI and checked against the existing entry
point ngp_nerf_network_inference on two calls in a row that share the static scratch array — the second with other inputs and a smaller
batch, so stale contents of the first would show.  Collected last on purpose (file name).
Passes alone in a fresh process and after the other tests (it touches no global state but the scratch array it owns)."""
import ctypes as C
import importlib

import numpy as np
import pytest
import torch

from .conftest import NERF_CFG_16x2

pytestmark = pytest.mark.gpu

MLP_TOL = 2e-3      # the tolerance of tests/test_ops_gpu.py for MLP outputs (the two paths run the same kernels: expected identical)


def _p(t):
    return C.c_void_p(t.data_ptr())


def test_inference_noalloc_matches_regular_entry_point_twice(ops):
    _lib = importlib.import_module("instant-ngp_b200._lib")
    lib = _lib.lib()
    cap = int(lib.ngp_noalloc_max_samples())
    assert cap == 16384
    net = ops.NerfNetwork(NERF_CFG_16x2)
    g = torch.Generator(device="cuda").manual_seed(5)
    nm = net.n_matrix_params
    params = torch.empty(net.n_params(), device="cuda")
    params[:nm] = (torch.rand(nm, device="cuda", generator=g) - 0.5) * 0.5
    params[nm:] = torch.rand(net.n_params() - nm, device="cuda", generator=g) - 0.5
    params = params.half()
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)

    def run_both(n, seed):
        gg = torch.Generator(device="cuda").manual_seed(seed)
        coords = torch.rand(n, 7, device="cuda", generator=gg)
        want = net.inference(params, coords)                                   # existing entry point (ngp_nerf_network_inference)
        got = torch.full((n, 4), float("nan"), dtype=torch.float16, device="cuda")
        _lib.check(lib.ngp_nerf_network_inference_noalloc(net._h, stream, _p(params), _p(coords), n, _p(got)))
        torch.cuda.synchronize()
        return got.float().cpu().numpy(), want.float().cpu().numpy()

    run_both(256, 1)                                                           # warm-up: module / kernel loading happens here
    free0 = torch.cuda.mem_get_info()[0]
    for n, seed in ((cap, 11), (1024, 12)):                                    # two calls in a row; full capacity, then a small batch
        gg = torch.Generator(device="cuda").manual_seed(seed)
        coords = torch.rand(n, 7, device="cuda", generator=gg)
        got = torch.full((n, 4), float("nan"), dtype=torch.float16, device="cuda")
        before = torch.cuda.mem_get_info()[0]
        _lib.check(lib.ngp_nerf_network_inference_noalloc(net._h, stream, _p(params), _p(coords), n, _p(got)))
        torch.cuda.synchronize()
        assert torch.cuda.mem_get_info()[0] == before, "the no-alloc entry point changed the device's free memory"
        want = net.inference(params, coords).float().cpu().numpy()
        got = got.float().cpu().numpy()
        assert np.isfinite(got).all()
        err = np.abs(got - want) / np.maximum(np.abs(want), 1.0)
        assert err.max() <= MLP_TOL, (n, err.max())
    assert free0 > 0

    # limits: beyond the static capacity / not a whole number of tiles -> error status + message, nothing launched
    coords = torch.rand(cap + 256, 7, device="cuda")
    out = torch.empty(cap + 256, 4, dtype=torch.float16, device="cuda")
    assert lib.ngp_nerf_network_inference_noalloc(net._h, stream, _p(params), _p(coords), cap + 256, _p(out)) != 0
    assert b"exceeds the static scratch capacity" in lib.ngp_last_error()
    assert lib.ngp_nerf_network_inference_noalloc(net._h, stream, _p(params), _p(coords), 100, _p(out)) != 0
    assert b"multiple of 128" in lib.ngp_last_error()
    assert lib.ngp_nerf_network_inference_noalloc(net._h, stream, _p(params), _p(coords), 0, _p(out)) == 0
