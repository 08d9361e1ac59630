"""Host-side snapshot codec (csrc/msgpack_mini.h, json_mini.h) against an independent implementation: the `msgpack` Python
package and the standard json module.  No device needed (the .ingp/.msgpack container of testbed.cu:4873-4936, :246-273)."""
import ctypes as C
import importlib
import json

import msgpack
import numpy as np
import pytest


@pytest.fixture(scope="module")
def L(ngp):
    return importlib.import_module("instant-ngp_b200._lib").lib()


def _roundtrip(L, data: bytes) -> bytes:
    cap = len(data) * 2 + 64
    out = C.create_string_buffer(cap); n = C.c_uint64()
    assert L.ngp_msgpack_roundtrip(data, len(data), out, cap, C.byref(n)) == 0, L.ngp_last_error()
    return out.raw[:n.value]


def test_msgpack_roundtrip_all_widths(L):
    rng = np.random.default_rng(0)
    obj = {
        "ints": [0, 1, 127, 128, 255, 256, 65535, 65536, 2**32 - 1, 2**32, 2**53, -1, -32, -33, -128, -129, -32768, -32769, -2**31, -2**31 - 1, -2**53],
        "floats": [0.5, -1.25, 1e-15, 0.33, 3.141592653589793, 1e300, float(np.float32(0.95))],
        "bools": [True, False, None],
        "strs": ["", "a" * 31, "b" * 32, "c" * 255, "d" * 256, "e" * 70000, "üñí"],
        "bins": [b"", bytes(rng.integers(0, 256, 255, dtype=np.uint8)), bytes(rng.integers(0, 256, 256, dtype=np.uint8)),
                 bytes(rng.integers(0, 256, 70000, dtype=np.uint8))],
        "arr15": list(range(15)), "arr16": list(range(16)), "arr70k": list(range(70000)),
        "map16": {f"k{i}": i for i in range(16)}, "map70k": {f"k{i}": i for i in range(70000)},
        "nested": {"encoding": {"otype": "HashGrid", "n_levels": 16, "per_level_scale": 1.3819}, "snapshot": {"params_binary": b"\x00\x3c" * 1000}},
    }
    data = msgpack.packb(obj, use_bin_type=True)
    back = _roundtrip(L, data)
    assert msgpack.unpackb(back, raw=False, strict_map_key=False) == obj
    # integers take their smallest encoding and exact float32 values are written as float32 (nlohmann::json::to_msgpack's rules)
    assert _roundtrip(L, msgpack.packb(200)) == b"\xcc\xc8"
    assert _roundtrip(L, msgpack.packb(-33)) == b"\xd0\xdf"
    assert _roundtrip(L, msgpack.packb(0.5)) == b"\xca\x3f\x00\x00\x00"
    assert _roundtrip(L, msgpack.packb(0.1)) == msgpack.packb(0.1)
    assert _roundtrip(L, back) == back                       # idempotent


def test_msgpack_rejects_malformed_input(L):
    good = msgpack.packb({"a": [1, 2, {"b": b"xyz"}]}, use_bin_type=True)
    out = C.create_string_buffer(256); n = C.c_uint64()
    for bad in (good[:-1], good + b"\x00", b"\xc1", b"\xdd\xff\xff\xff\xff", b"\x81\x01\x01"):
        assert L.ngp_msgpack_roundtrip(bad, len(bad), out, 256, C.byref(n)) != 0
        assert b"msgpack" in L.ngp_last_error()
    assert L.ngp_msgpack_roundtrip(good, len(good), out, 2, C.byref(n)) != 0 and n.value == len(good)     # too small: size is reported


def test_json_config_to_msgpack_matches_python(L):
    text = """{ // comments are allowed in the reference's configs
        "loss": {"otype": "Huber"},
        "optimizer": {"otype": "Ema", "decay": 0.95, "nested": {"otype": "ExponentialDecay", "decay_start": 20000, "decay_interval": 10000,
            "decay_base": 0.33, "nested": {"otype": "Adam", "learning_rate": 1e-2, "beta1": 0.9, "beta2": 0.99, "epsilon": 1e-15, "l2_reg": 1e-6}}},
        /* block comment */
        "encoding": {"otype": "HashGrid", "n_levels": 16, "n_features_per_level": 2, "log2_hashmap_size": 19, "base_resolution": 16},
        "flags": [true, false, null, -3, 2.0]
    }"""
    out = C.create_string_buffer(4096); n = C.c_uint64()
    assert L.ngp_json_to_msgpack(text.encode(), out, 4096, C.byref(n)) == 0, L.ngp_last_error()
    got = msgpack.unpackb(out.raw[:n.value], raw=False)
    stripped = "\n".join(l.split("//")[0] for l in text.split("\n")).replace("/* block comment */", "")
    want = json.loads(stripped)
    assert got == want
    assert isinstance(got["encoding"]["n_levels"], int) and isinstance(got["flags"][4], float)   # "2.0" stays a float, "16" an integer
    assert L.ngp_json_to_msgpack(b"{\"a\": ", out, 4096, C.byref(n)) != 0
