"""The native host runtime (C++): utilities, text format, parsers, filters and whole clusters
inside one process (in-memory van) and over TCP."""
import glob
import os

import numpy as np
import pytest

from parameter_server_b200.runtime import host


def test_crc32c_rfc3720_vectors():
    assert host.crc32c(b"\x00" * 32) == 0x8A9136AA
    assert host.crc32c(b"\xff" * 32) == 0x62A8AB43
    assert host.crc32c(bytes(range(32))) == 0x46DD794E
    assert host.crc32c(bytes(range(31, -1, -1))) == 0x113FDB5C
    assert host.crc32c(b"123456789") == 0xE3069283
    # extend(crc(a), b) == crc(a+b)
    assert host.crc32c(b"world", init=host.crc32c(b"hello ")) == host.crc32c(b"hello world")


def test_murmur3_reference_values():
    assert host.murmur3_32(b"", 0) == 0
    assert host.murmur3_32(b"", 1) == 0x514E28B7
    assert host.murmur3_32(b"hello", 0) == 0x248BFA47
    assert host.murmur3_32(b"The quick brown fox jumps over the lazy dog", 0) == 0x2E4FF723
    h = host.murmur3_128(b"The quick brown fox jumps over the lazy dog", 0)
    assert h == (0xE34BBC7BBC071B6C, 0x7A433CA9C49A9347)


@pytest.mark.parametrize("n", [0, 1, 7, 100, 70000, 300000])
def test_snappy_roundtrip(n):
    rng = np.random.RandomState(n)
    data = (rng.randint(0, 4, n).astype(np.uint8).tobytes() + b"abcabcabc" * (n // 50))
    c = host.snappy_compress(data)
    assert host.snappy_uncompress(c) == data
    if n >= 70000:
        assert len(c) < len(data)


def test_snappy_decodes_handwritten_stream():
    # varint length 11, literal "hello " (tag (6-1)<<2), copy offset 6 len 5 (type 1: 1|(5-4)<<2, 6)
    stream = bytes([11, 5 << 2]) + b"hello " + bytes([1 | (1 << 2), 6])
    assert host.snappy_uncompress(stream) == b"hello hello"


def test_range_even_divide_matches_python():
    from parameter_server_b200.utils.range import Range
    for b, e, n in [(0, 2**64 - 1, 3), (3, 1003, 7), (10, 47246, 297), (5, 6, 4)]:
        for i in range(n):
            r = Range(b, e).even_divide(n, i)
            assert host.range_even_divide(b, e, n, i) == (r.begin, r.end)


def test_textproto_features():
    t = """
    # comment
    a: 1 b { c: 'x\\ty' "z" [PS.LM.d]: 2.5 } e: [1, 2]
    f < g: true >
    """
    n = host.textproto_normalize(t, True)
    assert n == 'a: 1 b { c: "x\\tyz" [PS.LM.d]: 2.5 } e: 1 e: 2 f { g: true }'
    assert host.textproto_get(t, "b.[PS.LM.d]") == "2.5"
    assert host.textproto_get(t, "e[1]") == "2"
    with pytest.raises(Exception):
        host.textproto_normalize("a { b: 1", True)


def test_all_example_configs_parse():
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    files = sorted(glob.glob(os.path.join(root, "example", "linear", "*", "*.conf")))
    files += sorted(glob.glob("/root/reference/example/linear/*/*.conf"))
    files += sorted(glob.glob("/root/reference/docker/client/config/*.conf"))
    assert len(files) >= 6
    for f in files:
        text = open(f).read()
        assert host.textproto_get(text, "linear_method"), f
    ctr = open(os.path.join(root, "example/linear/ctr/batch_l1lr.conf")).read()
    assert host.textproto_get(ctr, "linear_method.darlin.max_block_delay") == "8"
    assert host.textproto_get(ctr, "linear_method.darlin.prior_fea_group[1]") == "120"
    assert host.textproto_get(ctr, "linear_method.darlin.[PS.LM.delta_max_value]") == "5"


def test_parsers_all_formats():
    ex, bad = host.parse_text("+1 3:0.5 10:2 77:1e-1\n-1 1:1\nbroken 3:1\n1 5:1 2:1\n", "libsvm")
    assert bad == 2 and len(ex) == 2
    assert ex[0][0][0] == 0 and ex[0][0][2].tolist() == [1.0]
    assert ex[0][1][1].tolist() == [3, 10, 77] and np.allclose(ex[0][1][2], [0.5, 2, 0.1])
    ex, bad = host.parse_text("0; 3 11 12 13; 5 7;\n1; 3 1;\n", "sparse_binary")
    assert bad == 0 and ex[0][0][2].tolist() == [-1.0]
    assert [s[0] for s in ex[0]] == [0, 3, 5] and ex[0][1][1].tolist() == [11, 12, 13]
    ex, _ = host.parse_text("1; 2 4:0.5 9:1.5;\n", "sparse")
    assert ex[0][1][1].tolist() == [4, 9] and ex[0][1][2].tolist() == [0.5, 1.5]
    ex, _ = host.parse_text("1; 1 0.5 0.25 4;\n", "dense")
    assert ex[0][1][2].tolist() == [0.5, 0.25, 4.0] and ex[0][1][1].size == 0
    ex, _ = host.parse_text("77 1 1 100:4 101:4 7:2\n", "adfea")
    assert [s[0] for s in ex[0]] == [0, 4, 2] and ex[0][1][1].tolist() == [100, 101]
    k = (5 << 54) | 123
    ex, _ = host.parse_text(f"0 99 | {k} {(5 << 54) | 7} {(9 << 54) | 1}\n", "terafea")
    assert [s[0] for s in ex[0]] == [0, 5, 9] and ex[0][1][1].tolist() == [k, (5 << 54) | 7]
    ex, _ = host.parse_text("1; 3 11; 5 7;\n", "sparse_binary", ignore_feature_group=True)
    assert [s[0] for s in ex[0]] == [0, 1] and ex[0][1][1].tolist() == [11, 7]


def test_example_binary_codec_matches_protobuf_wire_format():
    blob = host.example_roundtrip("+1 3:0.5 300:2", "libsvm")
    # Example{slot{id:0 val:[1]} slot{id:1 key:[3,300] val:[.5,2]}} in protobuf wire format
    import struct
    s0 = bytes([0x08, 0x00, 0x1A, 0x04]) + struct.pack("<f", 1.0)
    s1 = bytes([0x08, 0x01, 0x12, 0x03, 0x03, 0xAC, 0x02, 0x1A, 0x08]) + struct.pack("<ff", 0.5, 2.0)
    assert blob == bytes([0x0A, len(s0)]) + s0 + bytes([0x0A, len(s1)]) + s1


def test_countmin_never_underestimates_and_filters():
    rng = np.random.RandomState(0)
    keys = np.unique(rng.randint(0, 2**62, 5000).astype(np.uint64))
    counts = rng.randint(1, 9, keys.size).astype(np.uint8)
    keep, table = host.countmin_filter(keys, counts, n=20000, k=2, freq=4)
    assert keep[counts > 4].all()              # sketch estimates are upper bounds
    assert (~keep[counts <= 4]).mean() > 0.7   # most rare keys are dropped at this load
    assert table.max() <= 254


def test_bloom_filters():
    f_std = host.bloom_false_positive_rate(20000, 50000, 20000 * 10, 4, blocked=False)
    f_blk = host.bloom_false_positive_rate(20000, 50000, 20000 * 10, 4, blocked=True)
    assert f_std < 0.03 and f_blk < 0.08


@pytest.mark.parametrize("nbytes", [1, 2, 3])
def test_fixing_float_is_unbiased_and_bounded(nbytes):
    rng = np.random.RandomState(1)
    x = rng.uniform(-90, 90, 20000)
    x[:3] = [100, 0.1, -100]  # out of range values are clamped
    y, code = host.fixing_float_roundtrip(x, nbytes, -90, 90, seed=7)
    step = 180.0 / (2 ** (8 * nbytes) - 2)
    inside = np.abs(x) <= 90
    assert np.abs(y - x)[inside].max() <= step * 1.0001
    assert abs((y - x)[inside].mean()) < step * 0.05   # stochastic rounding: no systematic bias
    assert y[0] == pytest.approx(90, abs=step) and y[2] == pytest.approx(-90, abs=step)
    assert code.size == x.size * nbytes


def test_auc_accuracy_against_sklearn():
    from sklearn.metrics import roc_auc_score
    rng = np.random.RandomState(2)
    y = np.where(rng.rand(3000) > 0.4, 1.0, -1.0)
    p = y * 0.8 + rng.randn(3000)
    r = host.auc_accuracy(y, p)
    assert r["auc"] == pytest.approx(roc_auc_score(y > 0, p), abs=1e-9)
    assert r["bucketed_auc"] == pytest.approx(r["auc"], abs=1e-3)
    assert r["accuracy"] == pytest.approx(((p > 0) == (y > 0)).mean(), abs=1e-12)


def test_kv_cluster_inproc_and_tcp():
    # 3 servers hold keys 0..5 = 0.1*k, both workers push +1 to their 4 keys, worker 0 pulls
    want = np.array([0, 0.2, 0.4, 0.5], dtype=np.float32) + 2
    got = host.run_kv_cluster(3, 2, 6, [0, 2, 4, 5])
    assert np.allclose(got, want)
    got = host.run_kv_cluster(3, 2, 6, [0, 2, 4, 5], tcp=True, base_port=19200 + os.getpid() % 500)
    assert np.allclose(got, want)


@pytest.mark.parametrize("tcp", [False, True])
def test_van_ring(tcp):
    """ref src/test/van_test.cc: a message survives n laps around a ring of vans unchanged"""
    sent = host.van_ring(4, tcp=tcp, base_port=21000 + os.getpid() % 3000, payload=5000)
    assert sent >= 4 * 5000 * (8 + 4)          # 4 laps of keys + values leave node 0


def test_reassign_server_key_range():
    """ref src/test/reassign_server_key_range.cc: ManageNode::UPDATE re-partitions [0, 100)"""
    from parameter_server_b200.utils.range import Range
    got = host.run_reassign_scenario(3, 3, 0, 100)
    assert sorted(got) == ["S0", "S1", "S2", "W0", "W1", "W2"]
    for s in range(3):
        r = Range(0, 100).even_divide(3, s)
        assert got[f"S{s}"] == (r.begin, r.end)
    assert all(got[f"W{w}"] == (0, 100) for w in range(3))


@pytest.mark.parametrize("area", ["bitmap", "sarray", "matrix", "localizer", "match", "concurrent"])
def test_native_util_checks(area):
    """C++-side checks in capi.cc mirroring the reference's bitmap / shared_array / sparse_matrix /
    localizer / common unit tests"""
    assert host.selftest(area) >= 3
    with pytest.raises(Exception):
        host.selftest("no-such-area")


def test_kvmap_operators_apply_in_order():
    """KVMap (ref src/parameter/kv_map.h): PLUS, TIMES, MINUS, DIVIDE pushed back to back without
    waiting are applied in timestamp order on every shard"""
    np.testing.assert_allclose(host.run_kvmap_scenario(16), 7.0)
