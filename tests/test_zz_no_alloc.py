"""psb_darlin_evaluate_noalloc: the progress evaluation for callers that cannot allocate device
memory (all HBM already committed to the job's own pools, or a thread that must not block on the
allocator).  The entry point is called straight through the C ABI, as such a caller would, and
must reproduce psb_darlin_evaluate (caller-provided scratch) on calls in a row: each later
call sees the scratch array exactly as the one before left it (also with a larger grid after a
smaller one, where an old arrival counter lies inside the new range of partial sums).

Collected last on purpose (file name): if it fails under -x, every other test has already run.
"""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu


def _structs(n_rows, n_keys, seed):
    from parameter_server_b200.ops.darlin import DarlinCtrl, DarlinServer, DarlinWorker
    g = torch.Generator(device="cuda").manual_seed(seed)
    z = torch.randn(n_rows, generator=g, device="cuda", dtype=torch.float64) * 3
    w = torch.randn(n_keys, generator=g, device="cuda", dtype=torch.float64)
    w[torch.rand(n_keys, generator=g, device="cuda") < 0.5] = 0
    w[::17] = float("nan")                                    # KKT-filtered coordinates
    active = torch.randint(-2**31, 2**31 - 1, ((n_keys + 31) // 32,), generator=g, device="cuda",
                           dtype=torch.int32)
    rem = n_keys % 32
    if rem:
        active[-1] &= (1 << rem) - 1
    vio = torch.tensor([0.625], device="cuda", dtype=torch.float64)
    wk, sv = DarlinWorker(), DarlinServer()
    wk.z, wk.n_rows = z.data_ptr(), n_rows
    sv.w, sv.active, sv.violation, sv.n_keys = w.data_ptr(), active.data_ptr(), vio.data_ptr(), n_keys
    ctrl = DarlinCtrl(1e20, 0.75, 1.0, 5.0, 40, 0, 2, 0, 0, 0, 0)
    ctrl_dev = torch.frombuffer(bytearray(bytes(ctrl)), dtype=torch.uint8).cuda()
    return wk, sv, ctrl_dev, (z, w, active, vio)


def _read(buf):
    raw = buf.cpu()
    f = raw[:3].view(torch.float64).tolist()
    return dict(loss=f[0], penalty=f[1], violation=f[2], nnz_w=int(raw[3]), nnz_active=int(raw[4]),
                seq=int(raw[5]))


def test_noalloc_evaluate_matches_scratch_version_twice():
    from parameter_server_b200._lib import check, cuda_lib
    from parameter_server_b200.ops._sig import apply_cuda_sigs
    apply_cuda_sigs()
    lib = cuda_lib()
    lib.psb_darlin_evaluate_noalloc.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                                C.c_uint64, C.c_int32, C.c_void_p]
    lib.psb_darlin_evaluate_noalloc.restype = C.c_int
    lib.psb_darlin_evaluate_noalloc_max_grid.restype = C.c_int32
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    max_grid = lib.psb_darlin_evaluate_noalloc_max_grid()
    assert max_grid >= 148

    for call, (n_rows, n_keys, grid) in enumerate([(200_003, 47_236, 296), (31_111, 1_000, 64),
                                                 (150_001, 5_000, 512)]):  # smaller, then larger grid on the used array
        wk, sv, ctrl, keep = _structs(n_rows, n_keys, seed=call)
        ref_out = torch.zeros(6, dtype=torch.int64, device="cuda")
        new_out = torch.zeros(6, dtype=torch.int64, device="cuda")
        scratch = torch.zeros(4 * grid + 4, dtype=torch.float64, device="cuda")
        check(lib.psb_darlin_evaluate(C.byref(wk), C.byref(sv), ctrl.data_ptr(), ref_out.data_ptr(), None,
                                      0, 0, scratch.data_ptr(), grid, stream))
        check(lib.psb_darlin_evaluate_noalloc(C.byref(wk), C.byref(sv), ctrl.data_ptr(),
                                              new_out.data_ptr(), 0, grid, stream))
        torch.cuda.synchronize()
        ref, new = _read(ref_out), _read(new_out)
        assert new["nnz_w"] == ref["nnz_w"] and new["nnz_active"] == ref["nnz_active"]
        assert new["seq"] == ref["seq"] == 42
        for k in ("loss", "penalty", "violation"):
            assert abs(new[k] - ref[k]) <= 1e-9 * max(1.0, abs(ref[k])), (call, k, new[k], ref[k])
        # and against plain PyTorch
        z, w, active, vio = keep
        ok = ~torch.isnan(w) & (w != 0)
        assert abs(ref["loss"] - float(torch.log1p(torch.exp(-z)).sum())) <= 1e-9 * ref["loss"]
        assert ref["nnz_w"] == int(ok.sum())
        assert abs(ref["penalty"] - 0.75 * float(w[ok].abs().sum())) <= 1e-9 * ref["penalty"]

    # sizes beyond the fixed scratch are refused with a message, nothing is launched
    rc = lib.psb_darlin_evaluate_noalloc(C.byref(wk), C.byref(sv), ctrl.data_ptr(), new_out.data_ptr(), 0,
                                         max_grid + 1, stream)
    assert rc != 0 and b"exceeds the maximum" in lib.psb_last_error()
