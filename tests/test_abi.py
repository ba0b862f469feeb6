"""The C-ABI shared library loads without a GPU, exports every symbol include/b2train.h declares, and rejects bad
arguments with an error code + message instead of launching anything."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "b2train.h")


@pytest.fixture(scope="module")
def lib():
    from b200_train import _lib
    return _lib


def declared_symbols():
    text = open(HEADER).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    text = re.sub(r"#ifdef B2_BUILD_PROBES.*?#endif", "", text, flags=re.S)     # measurement aids: not in the shipped build
    return sorted(set(re.findall(r"\b(b2_[a-z0-9_]+)\s*\(", text)))


def test_header_declares_something():
    names = declared_symbols()
    assert "b2_adamw" in names and "b2_lmhead_ce_fwd" in names and "b2_peer_allreduce" in names
    assert len(names) >= 20


def test_every_declared_symbol_is_exported(lib):
    cdll = ctypes.CDLL(lib.LIB_PATH)
    missing = [n for n in declared_symbols() if not hasattr(cdll, n)]
    assert not missing, f"declared in b2train.h but not exported: {missing}"


def test_python_binding_covers_the_header(lib):
    assert sorted(lib.EXPORTS) == declared_symbols()


def test_version_and_sizes(lib):
    assert lib.lib.b2_version() == 100
    assert lib.lib.b2_sumsq_scratch_bytes() >= 4 * (32 + 148)
    assert lib.lib.b2_peer_flags_bytes() == 4 * 384      # barrier words, sumsq slots, per-block partials
    assert lib.lib.b2_lmhead_ce_workspace_bytes(8192) == 2 * 64 * 8192 * 8
    assert lib.lib.b2_lmhead_ce_workspace_bytes(1) == 2 * 64 * 128 * 8
    assert lib.lib.b2_gemm_workspace_bytes(100, 770) == 6 * 100 * 772 * 4


def test_bad_arguments_are_rejected_without_a_gpu(lib):
    L = lib.lib
    assert L.b2_adamw(None, None, None, None, 16, 0, 0, 0, 1e-3, 0.9, 0.999, 1e-8, 0.0, 1, None, None) == -1
    assert b"null" in L.b2_last_error()
    assert L.b2_adamw(16, 16, 16, 16, 16, 0, 0, 0, 1e-3, 0.9, 0.999, 1e-8, 0.0, 0, None, None) == -1
    assert b"step" in L.b2_last_error()
    assert L.b2_adamw(16, 16, 16, 16, 16, 7, 0, 0, 1e-3, 0.9, 0.999, 1e-8, 0.0, 1, None, None) == -1
    assert L.b2_adamw(8, 16, 16, 16, 16, 0, 0, 0, 1e-3, 0.9, 0.999, 1e-8, 0.0, 1, None, None) == -1
    assert b"aligned" in L.b2_last_error()
    assert L.b2_sumsq(16, 8, 3, 16, 16, 0, None) == -1
    assert L.b2_ce_rows_fwd(16, 0, 4, 0, 0, 16, 0, -100, 16, 16, 16, 16, None, None) == -1
    assert L.b2_lmhead_ce_fwd(16, 8, 16, 8, None, 16, 10, 8, 8, 3, -100, 1, 16, 0, 16, 16, 16, 16, None, None) == -1
    assert b"multiple of shift_T" in L.b2_last_error()
    assert L.b2_peer_allreduce(None, None, 9, 0, 0, 64, 1, 0.5, 1, 0, 0, 0, None, None) == -1
    assert L.b2_peer_allreduce(None, None, 2, 2, 0, 64, 1, 0.5, 1, 0, 0, 0, None, None) == -1
    with pytest.raises(lib.B2Error):
        lib.call("b2_scale", None, 8, 0, None, None)


def test_peer_slices_partition_the_buffer(lib):
    b, e = ctypes.c_size_t(), ctypes.c_size_t()
    for dtype, per_vec in ((0, 4), (1, 8)):
        for n in (per_vec, 64, 1 << 20, (1 << 20) + 64, 7 * 64):
            for world in (1, 2, 3, 4, 8):
                cover = []
                for r in range(world):
                    assert lib.lib.b2_peer_slice(n, dtype, world, r, ctypes.byref(b), ctypes.byref(e)) == 0
                    assert b.value % per_vec == 0 and b.value <= e.value <= n
                    cover.append((b.value, e.value))
                assert cover[0][0] == 0 and cover[-1][1] == n
                assert all(cover[i][1] == cover[i + 1][0] for i in range(world - 1))


def test_no_cpu_fallback(lib):
    import torch
    from b200_train import ops
    x = torch.zeros(64)
    with pytest.raises(lib.B2Error):
        ops.sumsq(x)
    with pytest.raises(lib.B2Error):
        ops.adamw_(x, x, x.clone(), x.clone(), lr=1e-3, step=1)


def test_new_f1_entry_points_reject_bad_arguments(lib):
    L = lib.lib
    # kept-logits forward: the logits buffer must exist, be 16-byte aligned, with a row pitch >= V that is a multiple of 8
    assert L.b2_lmhead_ce_fwd_keep(16, 8, 16, 8, None, 16, 10, 8, 8, 0, -100, 16, 0, 16, 16, 16, 16, None, None, 8, None) == -1
    assert b"null logits" in L.b2_last_error()
    assert L.b2_lmhead_ce_fwd_keep(16, 8, 16, 8, None, 16, 10, 8, 8, 0, -100, 16, 0, 16, 16, 16, 16, None, 16, 9, None) == -1
    assert b"ldl" in L.b2_last_error()
    assert L.b2_ce_dlogits_inplace(16, 12, 16, 0, 4, 10, 0, -100, 16, 16, None) == -1
    assert b"ldl" in L.b2_last_error()
    assert L.b2_ce_dlogits_inplace(None, 16, 16, 0, 4, 10, 0, -100, 16, 16, None) == -1
    assert L.b2_ce_dlogits_inplace(16, 16, 16, 0, 0, 10, 0, -100, 16, 16, None) == 0       # zero rows: nothing to do


def test_built_kernels_target_sm100a_use_the_tensor_path_and_do_not_spill():
    """Static check on the shipped library (cuobjdump, no GPU): every kernel is sm_100a code without local-memory
    spills, and the F1 kernels really contain the tcgen05 / TMA instructions (a silent fallback to mma.sync-era code
    would not)."""
    import shutil
    import subprocess
    from b200_train import _lib
    if shutil.which("cuobjdump") is None:
        pytest.skip("cuobjdump not on PATH")
    res = subprocess.run(["cuobjdump", "-res-usage", _lib.LIB_PATH], capture_output=True, text=True, check=True).stdout
    assert "sm_100a" in res and not re.search(r"arch = sm_(?!100a)", res)
    usage = re.findall(r"Function (\S+):\s*\n\s*REG:(\d+) STACK:(\d+) SHARED:(\d+) LOCAL:(\d+)", res)
    assert len(usage) >= 40
    assert all(int(local) == 0 for _, _, _, _, local in usage), [n for n, *_r, loc in usage if int(loc)]
    assert all(int(reg) <= 168 for n, reg, *_ in usage if "umma_kernel" in n)      # 384 threads x 168 = the register file
    sass = subprocess.run(["cuobjdump", "-sass", _lib.LIB_PATH], capture_output=True, text=True, check=True).stdout
    for mnemonic in ("UTCHMMA", "UTCBAR", "LDTM", "UTMALDG", "UTMASTG", "LDGMC"):
        assert mnemonic in sass, f"{mnemonic} missing from the built library"


def test_more_entry_points_reject_bad_arguments(lib):
    L = lib.lib
    assert L.b2_gemm_bf16(16, 0, 8, 16, 0, 8, 16, 1, 8, 0, 8, 8, 0, None, 0, None) == -1          # M = 0
    assert b"bad shape" in L.b2_last_error()
    assert L.b2_gemm_bf16(16, 0, 8, 16, 0, 8, 16, 5, 8, 8, 8, 8, 0, None, 0, None) == -1          # c_dtype
    assert L.b2_gemm_bf16(16, 0, 8, 16, 0, 8, 16, 1, 4, 8, 8, 8, 0, None, 0, None) == -1          # ldc < N
    assert L.b2_gather_windows(None, 10, None, 2, 4, None, None, None, None) == -1
    assert b"null" in L.b2_last_error()
    assert L.b2_gather_windows(16, 0, 16, 2, 4, 16, 16, None, None) == -1                          # no tokens
    assert L.b2_gather_windows(None, 0, None, 0, 4, None, None, None, None) == 0                   # empty batch: nothing to do
    assert L.b2_peer_allgather(None, None, 2, 0, 0, 12, 1, 1, 0, None) == -1                          # 12 bf16 != k * 16 bytes
    assert b"16 bytes" in L.b2_last_error()
    assert L.b2_peer_allgather(None, None, 2, 5, 0, 16, 1, 1, 0, None) == -1
    assert L.b2_symm_alloc(0, None, None) == -1
    assert L.b2_lmhead_ce_bwd_dlogits(16, 8, 16, 8, None, 16, 0, 4, 8, 10, 0, -100, 1, 16, 16, 16, 12, None) == -1   # ldd % 8
    assert b"ldd" in L.b2_last_error()


def test_without_a_gpu_device_calls_fail_loudly(lib):
    """No CPU path anywhere: where there is no device, anything that would touch one returns an error code and a message
    (this test only runs where CUDA is absent, e.g. the build container)."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    L = lib.lib
    p = ctypes.c_void_p()
    h = (ctypes.c_ubyte * 64)()
    rc = L.b2_symm_alloc(4096, ctypes.byref(p), h)
    assert rc == -2 and L.b2_last_error()                                   # B2_ERR_CUDA, with the driver's reason
    assert L.b2_device_count() <= 0


def test_shipped_library_has_no_measurement_aids(lib):
    """b2_debug_* entry points and the B2_DEBUG_* kernel branches exist only in a -DB2_BUILD_PROBES build."""
    cdll = ctypes.CDLL(lib.LIB_PATH)
    assert not hasattr(cdll, "b2_debug_peer_copy")
    assert "b2_debug_peer_copy" not in lib.EXPORTS
