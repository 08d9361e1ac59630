"""max_level_rand_training (Testbed::m_max_level_rand_training, python_api.cu:498): every training ray draws a level cut-off
2u (testbed_nerf.cu:1438,1775) that travels with its samples (:1652-1657,1915-1917) into GridEncoding::set_max_level_gpu (:3996,4004), which
zeroes the finer hash-grid levels of those rows in the forward pass and keeps their gradients away in the backward pass.  Operator level
(encoding forward / backward against the oracle) and one whole training step (rng stream, cut-offs per sample, network output, loss
gradients, compacted cut-offs, parameter gradients)."""
import importlib

import numpy as np
import pytest
import torch

from .test_nerf_gpu import BATCH, N_VIEWS, RES, SMALL_CFG
from .test_ops_gpu import ENC_TOL, GRID_CASES, _cuda, _grid

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("case", [GRID_CASES[0], GRID_CASES[2]])
@pytest.mark.parametrize("layout", [0, 1])
def test_encoding_with_per_row_max_level(ops, ref, case, layout):
    enc, cfg = _grid(ops, ref, case)
    D, L, F = case[0], case[1], case[2]
    n = 4096 + 40
    rng = np.random.default_rng(11)
    table = rng.uniform(-1, 1, enc.n_params()).astype(np.float16)
    x = ref.pcg32_floats(1337, 9, n * D).reshape(n, D)
    ml = (rng.random(n) * 2.0).astype(np.float32)               # what the sampler draws: half of the rows keep every level
    ml[:L + 1] = np.arange(L + 1, dtype=np.float32) / L           # the exact boundaries: level l is cut at max_level * L == l (within 1e-3)
    ml[L + 1] = 0.0; ml[L + 2] = 1.0; ml[L + 3] = 5.0
    C = enc.padded_output_width()
    enc.set_max_level_gpu(_cuda(ml))
    y = enc.forward(_cuda(table), _cuda(x), layout)
    if layout == 1:
        y = ops.blk8_to_rowmajor(y, (n + 7) // 8 * 8, C)[:n]
    y = y.float().cpu().numpy()
    yr = ref.grid_forward(cfg, table, x, C, max_level=ml).astype(np.float32)
    assert np.abs(y - yr).max() <= ENC_TOL * max(1.0, np.abs(yr).max())
    cut = (np.arange(L)[None, :] >= (ml[:, None] * np.float32(L) + np.float32(1e-3)))          # [n, L]: levels the row does not see
    assert cut.any() and (~cut).any()
    assert (y[:, :L * F].reshape(n, L, F)[cut] == 0).all() and (yr[:, :L * F].reshape(n, L, F)[cut] == 0).all()
    full = ref.grid_forward(cfg, table, x, C).astype(np.float32)
    assert np.array_equal(yr[:, :L * F].reshape(n, L, F)[~cut], full[:, :L * F].reshape(n, L, F)[~cut])   # the levels that stay are untouched

    # backward: rows scatter only into the levels they saw (the comparison is > instead of >= there)
    dy = np.zeros((n, C), np.float16); dy[:, :L * F] = (rng.normal(0, 1, (n, L * F)) / 8).astype(np.float16)
    dyd = _cuda(dy)
    if layout == 1:
        dyd = ops.rowmajor_to_blk8(_cuda(np.concatenate([dy, np.zeros(((n + 7) // 8 * 8 - n, C), np.float16)])))
    g = enc.backward(_cuda(x), dyd, None, layout).float().cpu().numpy().astype(np.float64)
    gr = ref.grid_backward(cfg, x, dy, enc.n_params(), max_level=ml)
    assert np.linalg.norm(g - gr) <= 1e-2 * np.linalg.norm(gr)
    g_all = ref.grid_backward(cfg, x, dy, enc.n_params())
    assert np.linalg.norm(g_all - gr) > 0.05 * np.linalg.norm(g_all)                             # the cut-off matters for this input
    enc.set_max_level_gpu(None)
    y2 = enc.forward(_cuda(table), _cuda(x), 0).float().cpu().numpy()
    assert np.abs(y2 - full).max() <= ENC_TOL * max(1.0, np.abs(full).max())                     # and it is gone again


@pytest.fixture(scope="module")
def env(ngp, ref):
    pyngp = importlib.import_module("instant-ngp_b200.pyngp")
    scene = importlib.import_module("instant-ngp_b200.scene")
    views = scene.make_views(N_VIEWS, RES, RES)
    images, keep = ref.images_from_views(views)
    return dict(pyngp=pyngp, scene=scene, views=views, images=images, keep=keep)


@pytest.mark.parametrize("lanes,n_warm", [(1, 0), (8, 3), (0, 37)])
def test_training_step_with_random_max_level(env, ref, lanes, n_warm):
    pyngp = env["pyngp"]
    tb = pyngp.Testbed(pyngp.TestbedMode.Nerf)
    tb.sampler_lanes_per_ray = lanes
    tb.set_nerf_dataset(env["views"], aabb_scale=1)
    tb.reload_network_from_json(SMALL_CFG)
    assert tb.max_level_rand_training is False
    tb.max_level_rand_training = True
    assert tb.max_level_rand_training is True
    for _ in range(n_warm):
        tb.train(BATCH)
    st = tb.stats
    rpb, mb, step = st.rays_per_batch, st.measured_batch_size_before_compaction, st.training_step
    max_inf = BATCH * 16 if mb == 0 else (min(mb, BATCH * 16) + 255) // 256 * 256
    params_before = tb.param_buffer(1).cpu().numpy().copy()                 # fp16 parameters the step computes with
    # the step in its two halves (the data-parallel entry points): the gradients exist between them, before the optimizer consumes and clears them
    import ctypes as C
    _lib = importlib.import_module("instant-ngp_b200._lib")
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    _lib.check(_lib.lib().ngp_testbed_train_begin(tb._h, stream, BATCH))
    torch.cuda.synchronize()
    g = tb.param_buffer(3).float().cpu().numpy().astype(np.float64)
    m, mb2 = C.c_uint32(), C.c_uint32()
    _lib.check(_lib.lib().ngp_testbed_pending_counters(tb._h, C.byref(m), C.byref(mb2)))
    _lib.check(_lib.lib().ngp_testbed_train_end(tb._h, stream, max(m.value, 1), max(mb2.value, 1)))

    aabb = ref.make_aabb((0, 0, 0), (1, 1, 1))
    bf = tb.density_grid(1).cpu().numpy()
    skip = 1 + step * (1 << 32)
    o = ref.generate_training_samples(rpb, aabb, max_inf, 1337, skip, env["images"], bf, 0, True, 0.0, max_level_rand=True)
    c = tb.debug_tensor("counters", torch.int32, 4).cpu().numpy().view(np.uint32)
    assert c[0] == o["ray_counter"] and c[1] == o["numsteps_counter"]
    nr = int(c[0]); n_written = int(o["numsteps"][:nr, 0].sum())
    assert nr > 0 and n_written > 0
    # the extra draw moves the ray's start offset: rays and samples are those of the shifted stream, not of the default one
    assert np.array_equal(tb.debug_tensor("rays", torch.float32, nr * 6).cpu().numpy().reshape(nr, 6), o["rays"][:nr])
    gcoords = tb.debug_tensor("coords", torch.float32, n_written * 7).cpu().numpy().reshape(n_written, 7)
    assert np.array_equal(gcoords, o["coords"][:n_written])
    plain = ref.generate_training_samples(rpb, aabb, max_inf, 1337, skip, env["images"], bf, 0, True, 0.0)
    assert not np.array_equal(plain["coords"][:min(n_written, 4096)], o["coords"][:min(n_written, 4096)])
    # per-sample cut-offs: 2u per ray, constant along the ray
    gml = tb.debug_tensor("max_level", torch.float32, n_written).cpu().numpy()
    assert np.array_equal(gml, o["max_level"][:n_written])
    assert 0.0 <= gml.min() and gml.max() < 2.0 and 0.3 < (gml >= 1.0).mean() < 0.7

    # the network saw the cut-offs: output of the inference pass against the oracle network on a slice of the samples
    rcfg = ref.nerf_cfg_from_json(SMALL_CFG)
    k = min(n_written, 3072)
    mlp_out = tb.debug_tensor("mlp_out", torch.float16, n_written * 4).cpu().numpy().reshape(n_written, 4)
    want = ref.nerf_forward(rcfg, params_before, o["coords"][:k], max_level=gml[:k]).astype(np.float32)
    got = mlp_out[:k].astype(np.float32)
    err = np.abs(got - want) / np.maximum(np.abs(want), 1.0)
    assert err.max() <= 4e-3, err.max()
    unmasked = ref.nerf_forward(rcfg, params_before, o["coords"][:k]).astype(np.float32)
    if n_warm:                                                              # (an untrained table is ~1e-4: masking it changes nothing visible)
        assert np.abs(unmasked - want).max() > 20 * np.abs(got - want).max()

    # loss kernel: one more draw before the background colour; the cut-offs of the kept samples
    mean_density = float(tb.debug_tensor("mean_density", torch.float32, 1).cpu().numpy()[0])
    cfg = ref.LossCfg(4, 2, 3, 1, 0, 1, (ref.f32 * 3)(0, 0, 0), 128.0, 0.1, mean_density)
    lo = ref.compute_loss(rpb, aabb, 1337, skip, BATCH, nr, cfg, env["images"], mlp_out, o["ray_indices"], o["rays"], o["numsteps"], o["coords"],
                          max_level_rand=True)
    assert c[2] == lo["compacted_counter"]
    nc = int(min(c[2], BATCH))
    gdl = tb.debug_tensor("dloss", torch.float16, BATCH * 4).cpu().numpy().reshape(BATCH, 4).astype(np.float32)
    rdl = lo["dloss"][:nc].astype(np.float32)
    scale = np.abs(rdl).max()
    assert np.abs(gdl[:nc] - rdl).max() <= 2e-3 * scale + 1e-7
    gmlc = tb.debug_tensor("max_level_compacted", torch.float32, BATCH).cpu().numpy()
    assert np.array_equal(gmlc[:nc], lo["max_level_compacted"][:nc])
    if nc < BATCH:                                                          # fill_rollover<float>(.., max_level_compacted) (:4066-4069)
        assert np.array_equal(gmlc[nc:], gmlc[np.arange(nc, BATCH) % nc])

    # parameter gradients of the step: the oracle network on the compacted batch with the same cut-offs
    gcc = tb.debug_tensor("coords_compacted", torch.float32, BATCH * 7).cpu().numpy().reshape(BATCH, 7)
    gdl16 = tb.debug_tensor("dloss", torch.float16, BATCH * 4).cpu().numpy().reshape(BATCH, 4)
    gr = ref.nerf_backward(rcfg, params_before, gcc, gdl16, max_level=gmlc)
    g_all = ref.nerf_backward(rcfg, params_before, gcc, gdl16)
    nm = int(ref.nerf_n_params(rcfg)[3])                                    # offset of the encoding = number of MLP weights
    gh, ghr, gha = g[nm:], gr[nm:], g_all[nm:]
    # (step 0: the table gradients are ~1e-6 per entry, i.e. fp16 subnormals on the device side: coarser quantisation)
    assert np.linalg.norm(gh - ghr) <= (2e-2 if n_warm else 5e-2) * np.linalg.norm(ghr), (np.linalg.norm(gh - ghr), np.linalg.norm(ghr))
    assert np.linalg.norm(gha - ghr) > 10 * np.linalg.norm(gh - ghr)        # without the cut-offs the table gradient is another one
    if n_warm:
        assert np.linalg.norm(g[:nm] - gr[:nm]) <= 2e-2 * np.linalg.norm(gr[:nm])


def test_max_level_rand_training_still_converges(env):
    pyngp = env["pyngp"]
    tb = pyngp.Testbed(pyngp.TestbedMode.Nerf)
    tb.set_nerf_dataset(env["views"], aabb_scale=1)
    tb.reload_network_from_json(SMALL_CFG)
    tb.max_level_rand_training = True
    for _ in range(17):
        tb.train(BATCH)
    first = tb.loss
    for _ in range(400):
        tb.train(BATCH)
    assert tb.loss < 0.5 * first, (first, tb.loss)
    tb.max_level_rand_training = False                                      # and back: the plain step again, on the same testbed
    for _ in range(32):
        tb.train(BATCH)
    assert np.isfinite(tb.loss)
