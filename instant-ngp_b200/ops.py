"""Operator-level Python mirrors of the reference's tcnn-style objects, over the C ABI.

Device memory, streams and (later) torch.distributed come from PyTorch — plumbing only; all arithmetic runs in
libngp_b200.so.  Names/arguments follow the reference surface (nerf_network.h, tcnn factories):
  Encoding(json, n_dims)            <-> tcnn::create_encoding
  NerfNetwork(json)                 <-> ngp::NerfNetwork
  NetworkWithInputEncoding(json, n_in, n_out)
  adam_step(...)                    <-> Trainer::optimizer_step
"""
import ctypes as C
import json as _json

import torch

from ._lib import AdamHyper, check, lib

LAYOUT_ROWMAJOR, LAYOUT_BLK8 = 0, 1
BATCH_GRANULARITY = 256


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _ptr(t):
    return C.c_void_p(t.data_ptr()) if t is not None else C.c_void_p(0)


def _cfg(j):
    return (j if isinstance(j, str) else _json.dumps(j)).encode()


def _dev(t, dtype):
    assert t.is_cuda and t.dtype == dtype and t.is_contiguous(), (t.device, t.dtype, t.is_contiguous())
    return t


def blk8_to_rowmajor(t, n, c):
    """[n*c] blk8 fp16 -> [n, c] row-major (torch view ops; test/debug helper)."""
    return t.view(n // 8, c // 8, 8, 8).permute(0, 2, 1, 3).reshape(n, c)


def rowmajor_to_blk8(t):
    n, c = t.shape
    return t.view(n // 8, 8, c // 8, 8).permute(0, 2, 1, 3).contiguous().view(-1)


class Encoding:
    def __init__(self, config, n_dims_to_encode):
        self._h = C.c_void_p()
        check(lib().ngp_encoding_create(_cfg(config), n_dims_to_encode, C.byref(self._h)))
        self.n_dims = n_dims_to_encode

    def __del__(self):
        if getattr(self, "_h", None) and lib is not None:     # `lib` is already torn down when the interpreter exits
            lib().ngp_encoding_destroy(self._h)
            self._h = None

    def n_params(self):
        return lib().ngp_encoding_n_params(self._h)

    def padded_output_width(self):
        return lib().ngp_encoding_padded_output_width(self._h)

    def level_table(self):
        n = lib().ngp_encoding_level_table(self._h, None, None, None)
        off = (C.c_uint32 * (n + 1))(); sc = (C.c_float * n)(); res = (C.c_uint32 * n)()
        lib().ngp_encoding_level_table(self._h, off, sc, res)
        return list(off), list(sc), list(res)

    def initialize_params(self, seed=1337):
        p = torch.empty(self.n_params(), dtype=torch.float32, device="cuda")
        check(lib().ngp_encoding_initialize_params(self._h, _stream(), seed, _ptr(p)))
        return p

    def set_max_level_gpu(self, max_level):
        """GridEncoding::set_max_level_gpu: a float32 CUDA tensor with one level cut-off per row for the calls that follow, or None"""
        if max_level is not None:
            _dev(max_level, torch.float32)
        self._max_level = max_level                     # keeps the tensor alive while the library holds its address
        check(lib().ngp_encoding_set_max_level_gpu(self._h, None if max_level is None else _ptr(max_level)))

    def forward(self, params, x, layout=LAYOUT_ROWMAJOR):
        _dev(params, torch.float16); _dev(x, torch.float32)
        n = x.shape[0]
        npad = (n + 7) // 8 * 8
        y = torch.zeros(npad * self.padded_output_width(), dtype=torch.float16, device="cuda")
        check(lib().ngp_encoding_forward(self._h, _stream(), _ptr(params), _ptr(x), x.shape[1], n, _ptr(y), layout))
        return y.view(npad, self.padded_output_width())[:n] if layout == LAYOUT_ROWMAJOR else y

    def backward(self, x, dy, grad=None, layout=LAYOUT_ROWMAJOR):
        _dev(x, torch.float32); _dev(dy, torch.float16)
        if grad is None:
            grad = torch.zeros(self.n_params(), dtype=torch.float16, device="cuda")
        check(lib().ngp_encoding_backward(self._h, _stream(), _ptr(x), x.shape[1], x.shape[0], _ptr(dy), layout, _ptr(grad)))
        return grad

    def backward_input(self, params, x, dy, layout=LAYOUT_ROWMAJOR):
        dx = torch.empty(x.shape[0], self.n_dims, dtype=torch.float32, device="cuda")
        check(lib().ngp_encoding_backward_input(self._h, _stream(), _ptr(params), _ptr(x), x.shape[1], x.shape[0], _ptr(dy), layout, _ptr(dx)))
        return dx

    def corners(self, x, level):
        n = x.shape[0]; k = 1 << self.n_dims
        idx = torch.empty(n, k, dtype=torch.int32, device="cuda"); w = torch.empty(n, k, dtype=torch.float32, device="cuda")
        check(lib().ngp_encoding_corners(self._h, _stream(), _ptr(x), x.shape[1], n, level, _ptr(idx), _ptr(w)))
        return idx, w


class NerfNetwork:
    def __init__(self, config):
        self._h = C.c_void_p()
        check(lib().ngp_nerf_network_create(_cfg(config), C.byref(self._h)))
        a, b, c, d = C.c_uint64(), C.c_uint64(), C.c_uint64(), C.c_uint64()
        check(lib().ngp_nerf_network_layout(self._h, C.byref(a), C.byref(b), C.byref(c), C.byref(d)))
        self.off_density, self.off_rgb, self.off_encoding, self.n_matrix_params = a.value, b.value, c.value, d.value

    def __del__(self):
        if getattr(self, "_h", None) and lib is not None:     # `lib` is already torn down when the interpreter exits
            lib().ngp_nerf_network_destroy(self._h)
            self._h = None

    def n_params(self):
        return lib().ngp_nerf_network_n_params(self._h)

    def initialize_params(self, seed=1337):
        p = torch.empty(self.n_params(), dtype=torch.float32, device="cuda")
        check(lib().ngp_nerf_network_initialize_params(self._h, _stream(), seed, _ptr(p)))
        return p

    def inference(self, params, coords):
        _dev(params, torch.float16); _dev(coords, torch.float32)
        n = coords.shape[0]
        out = torch.empty(n, 4, dtype=torch.float16, device="cuda")
        check(lib().ngp_nerf_network_inference(self._h, _stream(), _ptr(params), _ptr(coords), n, _ptr(out)))
        return out

    def density(self, params, pos):
        _dev(params, torch.float16); _dev(pos, torch.float32)
        n = pos.shape[0]
        out = torch.empty(n, dtype=torch.float16, device="cuda")
        check(lib().ngp_nerf_network_density(self._h, _stream(), _ptr(params), _ptr(pos), pos.shape[1], n, _ptr(out)))
        return out

    def forward_backward(self, params, coords, dL_dout, grads=None, input_gradients=False):
        """input_gradients: also return dL/d(coords) [n, 7] fp32 (position through the hash grid, direction through SH4; the dt column stays 0)"""
        _dev(params, torch.float16); _dev(coords, torch.float32); _dev(dL_dout, torch.float16)
        if grads is None:
            grads = torch.empty(self.n_params(), dtype=torch.float16, device="cuda")
        if not input_gradients:
            check(lib().ngp_nerf_network_forward_backward(self._h, _stream(), _ptr(params), _ptr(coords), coords.shape[0], _ptr(dL_dout), _ptr(grads)))
            return grads
        dc = torch.zeros(coords.shape[0], 7, dtype=torch.float32, device="cuda")
        check(lib().ngp_nerf_network_forward_backward_input(self._h, _stream(), _ptr(params), _ptr(coords), coords.shape[0], _ptr(dL_dout), _ptr(grads), _ptr(dc)))
        return grads, dc


class NetworkWithInputEncoding:
    def __init__(self, config, n_input_dims, n_output_dims):
        self._h = C.c_void_p()
        check(lib().ngp_encoded_mlp_create(_cfg(config), n_input_dims, n_output_dims, C.byref(self._h)))
        a, b, c = C.c_uint64(), C.c_uint64(), C.c_uint64()
        check(lib().ngp_encoded_mlp_layout(self._h, C.byref(a), C.byref(b), C.byref(c)))
        self.off_mlp, self.off_encoding, self.n_matrix_params = a.value, b.value, c.value

    def __del__(self):
        if getattr(self, "_h", None) and lib is not None:     # `lib` is already torn down when the interpreter exits
            lib().ngp_encoded_mlp_destroy(self._h)
            self._h = None

    def n_params(self):
        return lib().ngp_encoded_mlp_n_params(self._h)

    def initialize_params(self, seed=1337):
        p = torch.empty(self.n_params(), dtype=torch.float32, device="cuda")
        check(lib().ngp_encoded_mlp_initialize_params(self._h, _stream(), seed, _ptr(p)))
        return p

    def inference(self, params, x):
        n = x.shape[0]
        out = torch.empty(n, 16, dtype=torch.float16, device="cuda")
        check(lib().ngp_encoded_mlp_inference(self._h, _stream(), _ptr(params), _ptr(x), x.shape[1], n, _ptr(out)))
        return out

    def forward_backward(self, params, x, dL_dout, grads=None):
        if grads is None:
            grads = torch.empty(self.n_params(), dtype=torch.float16, device="cuda")
        check(lib().ngp_encoded_mlp_forward_backward(self._h, _stream(), _ptr(params), _ptr(x), x.shape[1], x.shape[0], _ptr(dL_dout), _ptr(grads)))
        return grads


def adam_step(master, params, grads, m, v, steps, *, learning_rate, beta1, beta2, epsilon, l2_reg, loss_scale, n_matrix_params,
              optimize_matrix_params=True, optimize_non_matrix_params=True, ema=None, ema_half=None, ema_decay=0.0, ema_step=1):
    h = AdamHyper(learning_rate, beta1, beta2, epsilon, l2_reg, loss_scale, n_matrix_params, int(optimize_matrix_params),
                  int(optimize_non_matrix_params), ema_decay, ema_step)
    check(lib().ngp_adam_step(_stream(), C.byref(h), master.numel(), _ptr(master), _ptr(params), _ptr(grads), _ptr(m), _ptr(v), _ptr(steps),
                              _ptr(ema), _ptr(ema_half)))
