"""ctypes bindings of the C++ host runtime (libpsb200_host.so, C ABI in csrc/host/capi.cc).

The host runtime is the CPU plumbing path: postoffice / van / executor / customers, KV
containers, filters, parsers.  These bindings expose the pieces Python needs: utilities for
tests and tools, the data parsers, and whole clusters run inside the current process.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from .._lib import hcheck, host_lib

TEXT_FORMATS = {"dense": 1, "sparse": 2, "sparse_binary": 3, "adfea": 4, "libsvm": 5, "terafea": 6}


def _lib():
    lib = host_lib()
    if not getattr(lib, "_psh_sigs", False):
        u8p, u64p, f32p, f64p, i64p = (C.POINTER(C.c_uint8), C.POINTER(C.c_uint64),
                                       C.POINTER(C.c_float), C.POINTER(C.c_double),
                                       C.POINTER(C.c_int64))
        lib.psh_crc32c.argtypes = [C.c_char_p, C.c_int64]
        lib.psh_crc32c.restype = C.c_uint32
        lib.psh_crc32c_extend.argtypes = [C.c_uint32, C.c_char_p, C.c_int64]
        lib.psh_crc32c_extend.restype = C.c_uint32
        lib.psh_murmur3_x86_32.argtypes = [C.c_char_p, C.c_int32, C.c_uint32]
        lib.psh_murmur3_x86_32.restype = C.c_uint32
        lib.psh_murmur3_x64_128.argtypes = [C.c_char_p, C.c_int32, C.c_uint32, u64p]
        lib.psh_sketch_hash.argtypes = [C.c_uint64]
        lib.psh_sketch_hash.restype = C.c_uint32
        lib.psh_snappy_max_compressed.argtypes = [C.c_int64]
        lib.psh_snappy_max_compressed.restype = C.c_int64
        lib.psh_snappy_compress.argtypes = [C.c_char_p, C.c_int64, C.c_char_p, i64p]
        lib.psh_snappy_uncompressed_length.argtypes = [C.c_char_p, C.c_int64, i64p]
        lib.psh_snappy_uncompress.argtypes = [C.c_char_p, C.c_int64, C.c_char_p]
        lib.psh_range_even_divide.argtypes = [C.c_uint64, C.c_uint64, C.c_uint64, C.c_uint64, u64p]
        lib.psh_textproto_normalize.argtypes = [C.c_char_p, C.c_int, C.c_char_p, C.c_int64, i64p]
        lib.psh_textproto_get.argtypes = [C.c_char_p, C.c_char_p, C.c_char_p, C.c_int64, i64p]
        lib.psh_parse_text.argtypes = [C.c_char_p, C.c_int32, C.c_int32, i64p, C.c_int64, u64p,
                                       C.c_int64, f32p, C.c_int64, i64p]
        lib.psh_example_roundtrip.argtypes = [C.c_char_p, C.c_int32, C.c_char_p, C.c_int64, i64p]
        lib.psh_slot_reader_load.argtypes = [C.c_char_p, C.c_char_p, C.c_char_p, C.c_int64, i64p]
        lib.psh_slot_reader_arrays.argtypes = [C.c_char_p, C.c_char_p, C.c_int32, u64p, C.c_int64,
                                               i64p, C.c_int64, f32p, C.c_int64, i64p]
        lib.psh_countmin_filter.argtypes = [u64p, u8p, C.c_int64, C.c_int32, C.c_int32, C.c_int32,
                                            u8p, u8p]
        lib.psh_bloom_fpr.argtypes = [C.c_int64, C.c_int64, C.c_int64, C.c_int32, C.c_int32, f64p]
        lib.psh_fixing_float_roundtrip.argtypes = [f64p, C.c_int64, C.c_int32, C.c_double,
                                                   C.c_double, C.c_uint64, f64p, u8p]
        lib.psh_auc_accuracy.argtypes = [f64p, f64p, C.c_int64, f64p]
        lib.psh_run_cluster.argtypes = [C.c_char_p, C.c_int32, C.c_int32, C.c_int32, C.c_int32,
                                        C.c_int32, C.c_char_p, C.c_int64, i64p]
        lib.psh_run_kv_cluster.argtypes = [C.c_int32, C.c_int32, C.c_int32, u64p, C.c_int32,
                                           C.c_int32, C.c_int32, f32p]
        lib._psh_sigs = True
    return lib


def _ptr(a: np.ndarray, ctype):
    return a.ctypes.data_as(C.POINTER(ctype))


def _str_out(fn, *args, cap: int = 1 << 22) -> str:
    buf = C.create_string_buffer(cap)
    n = C.c_int64(0)
    hcheck(fn(*args, buf, cap, C.byref(n)))
    return buf.raw[:n.value].decode()


def crc32c(data: bytes, init: int | None = None) -> int:
    lib = _lib()
    return lib.psh_crc32c(data, len(data)) if init is None else lib.psh_crc32c_extend(init, data, len(data))


def murmur3_32(data: bytes, seed: int = 0) -> int:
    return _lib().psh_murmur3_x86_32(data, len(data), seed)


def murmur3_128(data: bytes, seed: int = 0) -> tuple:
    out = (C.c_uint64 * 2)()
    _lib().psh_murmur3_x64_128(data, len(data), seed, out)
    return out[0], out[1]


def sketch_hash(key: int) -> int:
    return _lib().psh_sketch_hash(key)


def snappy_compress(data: bytes) -> bytes:
    lib = _lib()
    buf = C.create_string_buffer(lib.psh_snappy_max_compressed(len(data)))
    n = C.c_int64(0)
    hcheck(lib.psh_snappy_compress(data, len(data), buf, C.byref(n)))
    return buf.raw[:n.value]


def snappy_uncompress(data: bytes) -> bytes:
    lib = _lib()
    n = C.c_int64(0)
    hcheck(lib.psh_snappy_uncompressed_length(data, len(data), C.byref(n)))
    buf = C.create_string_buffer(max(n.value, 1))
    hcheck(lib.psh_snappy_uncompress(data, len(data), buf))
    return buf.raw[:n.value]


def range_even_divide(begin: int, end: int, n: int, i: int) -> tuple:
    out = (C.c_uint64 * 2)()
    _lib().psh_range_even_divide(begin, end, n, i, out)
    return out[0], out[1]


def textproto_normalize(text: str, single_line: bool = False) -> str:
    return _str_out(_lib().psh_textproto_normalize, text.encode(), int(single_line))


def textproto_get(text: str, path: str) -> str:
    return _str_out(_lib().psh_textproto_get, text.encode(), path.encode())


def parse_text(text: str, fmt: str, ignore_feature_group: bool = False):
    """Parse text lines with the native ExampleParser.

    Returns (examples, n_bad) where examples is a list of lists of (slot_id, keys, vals)."""
    lib = _lib()
    n_lines = text.count("\n") + 1
    n_tok = len(text.split()) + len(text) // 2 + 16
    meta = np.zeros(n_lines * (3 * 64 + 1) + 16, dtype=np.int64)
    keys = np.zeros(n_tok, dtype=np.uint64)
    vals = np.zeros(n_tok, dtype=np.float32)
    counts = np.zeros(4, dtype=np.int64)
    hcheck(lib.psh_parse_text(text.encode(), TEXT_FORMATS[fmt], int(ignore_feature_group),
                              _ptr(meta, C.c_int64), meta.size, _ptr(keys, C.c_uint64), keys.size,
                              _ptr(vals, C.c_float), vals.size, _ptr(counts, C.c_int64)))
    out, m, k, v = [], 0, 0, 0
    for _ in range(int(counts[0])):
        ns = int(meta[m]); m += 1
        ex = []
        for _ in range(ns):
            sid, nk, nv = int(meta[m]), int(meta[m + 1]), int(meta[m + 2]); m += 3
            ex.append((sid, keys[k:k + nk].copy(), vals[v:v + nv].copy()))
            k += nk; v += nv
        out.append(ex)
    return out, int(counts[1])


def example_roundtrip(line: str, fmt: str) -> bytes:
    buf = C.create_string_buffer(1 << 20)
    n = C.c_int64(0)
    hcheck(_lib().psh_example_roundtrip(line.encode(), TEXT_FORMATS[fmt], buf, 1 << 20, C.byref(n)))
    return buf.raw[:n.value]


def slot_reader_info(data_conf: str, cache_prefix: str) -> str:
    return _str_out(_lib().psh_slot_reader_load, data_conf.encode(), cache_prefix.encode())


def slot_reader_arrays(data_conf: str, cache_prefix: str, slot: int, cap: int = 1 << 24):
    idx = np.zeros(cap, dtype=np.uint64)
    off = np.zeros(cap, dtype=np.int64)
    val = np.zeros(cap, dtype=np.float32)
    counts = np.zeros(3, dtype=np.int64)
    hcheck(_lib().psh_slot_reader_arrays(data_conf.encode(), cache_prefix.encode(), slot,
                                         _ptr(idx, C.c_uint64), cap, _ptr(off, C.c_int64), cap,
                                         _ptr(val, C.c_float), cap, _ptr(counts, C.c_int64)))
    return idx[:counts[0]].copy(), off[:counts[1]].copy(), val[:counts[2]].copy()


def countmin_filter(keys: np.ndarray, counts: np.ndarray, n: int, k: int, freq: int):
    keys = np.ascontiguousarray(keys, dtype=np.uint64)
    counts = np.ascontiguousarray(counts, dtype=np.uint8)
    keep = np.zeros(keys.size, dtype=np.uint8)
    table = np.zeros(max(n, 64), dtype=np.uint8)
    hcheck(_lib().psh_countmin_filter(_ptr(keys, C.c_uint64), _ptr(counts, C.c_uint8), keys.size, n, k,
                                      freq, _ptr(keep, C.c_uint8), _ptr(table, C.c_uint8)))
    return keep.astype(bool), table


def bloom_false_positive_rate(n_insert: int, n_probe: int, bits: int, k: int, blocked: bool) -> float:
    out = C.c_double(0)
    hcheck(_lib().psh_bloom_fpr(n_insert, n_probe, bits, k, int(blocked), C.byref(out)))
    return out.value


def fixing_float_roundtrip(x: np.ndarray, nbytes: int, lo: float, hi: float, seed: int = 1):
    x = np.ascontiguousarray(x, dtype=np.float64)
    out = np.zeros_like(x)
    code = np.zeros(x.size * nbytes, dtype=np.uint8)
    hcheck(_lib().psh_fixing_float_roundtrip(_ptr(x, C.c_double), x.size, nbytes, lo, hi, seed,
                                             _ptr(out, C.c_double), _ptr(code, C.c_uint8)))
    return out, code


def auc_accuracy(label: np.ndarray, predict: np.ndarray):
    label = np.ascontiguousarray(label, dtype=np.float64)
    predict = np.ascontiguousarray(predict, dtype=np.float64)
    out = np.zeros(3)
    hcheck(_lib().psh_auc_accuracy(_ptr(label, C.c_double), _ptr(predict, C.c_double), label.size,
                                   _ptr(out, C.c_double)))
    return dict(auc=out[0], accuracy=out[1], bucketed_auc=out[2])


def run_cluster(app_conf: str, num_servers: int, num_workers: int, tcp: bool = False,
                base_port: int = 9100, message_compression: bool = True) -> str:
    """1 scheduler + servers + workers inside this process (threads); returns the report text."""
    return _str_out(_lib().psh_run_cluster, app_conf.encode(), num_servers, num_workers, int(tcp),
                    base_port, int(message_compression))


def run_kv_cluster(num_servers: int, num_workers: int, n_keys: int, pull_keys, tcp: bool = False,
                   base_port: int = 9200) -> np.ndarray:
    keys = np.ascontiguousarray(pull_keys, dtype=np.uint64)
    out = np.zeros(keys.size, dtype=np.float32)
    hcheck(_lib().psh_run_kv_cluster(num_servers, num_workers, n_keys, _ptr(keys, C.c_uint64), keys.size,
                                     int(tcp), base_port, _ptr(out, C.c_float)))
    return out


def run_replace_scenario(n_keys: int):
    """Kill server S1 of a 2-server in-process cluster with replication and swap the spare U0 in
    (ManageNode::REPLACE). Returns (values the worker pulls afterwards, S1's key range)."""
    out = np.zeros(n_keys, dtype=np.float32)
    rng = (C.c_uint64 * 2)()
    hcheck(_lib().psh_run_replace_scenario(n_keys, _ptr(out, C.c_float), rng))
    return out, (int(rng[0]), int(rng[1]))


def run_heartbeat_scenario(interval_s: int = 1, run_ms: int = 2600, dead_timeout_s: int = 60):
    """Returns (dashboard table of the scheduler, number of nodes overdue by dead_timeout_s)."""
    buf = C.create_string_buffer(1 << 16)
    n = C.c_int64(0)
    dead = C.c_int32(0)
    hcheck(_lib().psh_run_heartbeat_scenario(interval_s, run_ms, dead_timeout_s, buf, 1 << 16, C.byref(n),
                                             C.byref(dead)))
    return buf.raw[:n.value].decode(), int(dead.value)


def van_ring(n: int = 3, tcp: bool = True, base_port: int = 9600, payload: int = 1000) -> int:
    """A message with `payload` keys and values travels n laps around a ring of n vans
    (ref src/test/van_test.cc); returns the bytes node 0 sent."""
    b = C.c_int64(0)
    hcheck(_lib().psh_van_ring(n, int(tcp), base_port, payload, C.byref(b)))
    return int(b.value)


def run_reassign_scenario(num_servers: int, num_workers: int, lo: int, hi: int) -> dict:
    """{node id: (begin, end)} after the scheduler re-partitioned [lo, hi) over the servers and
    broadcast ManageNode::UPDATE (ref src/test/reassign_server_key_range.cc)."""
    text = _str_out(_lib().psh_run_reassign_scenario, num_servers, num_workers, C.c_uint64(lo), C.c_uint64(hi))
    out = {}
    for line in text.strip().splitlines():
        i, b, e = line.split()
        out[i] = (int(b), int(e))
    return out


def selftest(area: str = "all") -> int:
    """Run the native unit checks of one area (bitmap, sarray, matrix, localizer, match,
    concurrent) or all of them; returns the number of checks, raises on the first failure."""
    n = C.c_int32(0)
    hcheck(_lib().psh_selftest(area.encode(), C.byref(n)))
    return int(n.value)


def run_kvmap_scenario(n_keys: int = 16) -> np.ndarray:
    """KVMap on 2 servers: one worker pushes +5, *3, -1, /2 on every key and pulls (expects 7)."""
    out = np.zeros(n_keys, dtype=np.float64)
    hcheck(_lib().psh_run_kvmap_scenario(n_keys, _ptr(out, C.c_double)))
    return out
