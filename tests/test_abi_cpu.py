"""CPU-side checks of the boundary: the shared library loads, exports every
symbol include/raft_corr.h declares, its host-only entry points work, and the
Python drop-in refuses to run without a GPU (no fallback)."""
import ctypes as C
import os
import re

import pytest
import torch

from raft_stir_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "raft_corr.h")).read()
    return sorted(set(re.findall(r"RC_API[^;]*?\b(rc_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    names = declared_symbols()
    assert len(names) >= 14
    handle = C.CDLL(_lib.LIB_PATH)
    for n in names:
        assert hasattr(handle, n), f"{n} declared in raft_corr.h but not exported"
    assert set(names) == set(_lib.SIGNATURES), "ctypes binding and header disagree"


def test_abi_version_and_layout():
    L = _lib.lib()
    assert L.rc_abi_version() == _lib.RC_ABI_VERSION
    lay = _lib.pyramid_layout(1, 55, 128, 4, _lib.RC_VOL_BF16)
    assert list(lay.H)[:4] == [55, 27, 13, 6] and list(lay.W)[:4] == [128, 64, 32, 16]
    assert lay.N == 7040 and lay.elem_bytes == 2
    # 64-byte tiles of 4 rows x 8 bf16: 55x128 -> 14x16 tiles, 27x64 -> 7x8, 13x32 -> 4x4, 6x16 -> 2x2
    assert (lay.tile_h, lay.tile_w) == (4, 8)
    assert list(lay.tiles_y)[:4] == [14, 7, 4, 2] and list(lay.tiles_x)[:4] == [16, 8, 4, 2]
    # SURVEY 8 counts 65 331 200 unpadded elements per Sintel pair; tile padding adds 3.4 %
    assert lay.total_elems == 7040 * 32 * (14 * 16 + 7 * 8 + 4 * 4 + 2 * 2)
    lay = _lib.pyramid_layout(10, 46, 62, 4, _lib.RC_VOL_F32)
    assert (lay.tile_h, lay.tile_w) == (4, 4)
    assert list(lay.W)[:4] == [62, 31, 15, 7] and list(lay.tiles_x)[:4] == [16, 8, 4, 2]
    for l in range(4):
        assert lay.slab[l] == lay.tiles_y[l] * lay.tiles_x[l] * 16 and lay.offset[l] % 64 == 0


def test_layout_rejects_bad_arguments():
    with pytest.raises(RuntimeError, match="no level"):
        _lib.pyramid_layout(1, 8, 8, 5, 0)     # the reference would divide by zero / pool to nothing
    with pytest.raises(RuntimeError):
        _lib.pyramid_layout(1, 16, 16, 7, 0)
    with pytest.raises(RuntimeError):
        _lib.pyramid_layout(1, 16, 16, 4, 9)


def test_workspace_sizes():
    L = _lib.lib()
    assert L.rc_build_workspace_bytes(1, 256, 55, 128, _lib.RC_MATH_F32) == 0
    # bf16 repack: N*D*2 for fmap1 + 28 blocks * 256 rows * D * 2 for fmap2
    assert L.rc_build_workspace_bytes(1, 256, 55, 128, _lib.RC_MATH_BF16) == 7040 * 256 * 2 + 28 * 256 * 256 * 2
    assert L.rc_build_bwd_workspace_bytes(1, 256, 55, 128, 4) > 0


def test_tensor_core_build_rejects_unsupported_sizes_before_launching():
    """D % 8 != 0 or more than 4 levels with a tensor-core math mode: RC_EINVAL from the argument checks, i.e. before
    the pack kernel could be queued (no CUDA call has happened: this runs without a GPU; the pointers are never
    dereferenced)."""
    L = _lib.lib()
    fake = 0x1000   # non-null, 256-byte aligned
    for D, levels in ((3, 4), (12, 2), (64, 5)):
        for math in (_lib.RC_MATH_BF16, _lib.RC_MATH_F16, _lib.RC_MATH_TF32):
            r = L.rc_build_pyramid(fake, fake, 1, D, 64, 64, levels, _lib.RC_VOL_BF16, math, fake, fake, 1 << 40, None)
            assert r == -1, (D, levels, math, r)
            assert b"tensor-core build" in L.rc_last_error()
        r = L.rc_build_pyramid_static(fake, fake, 1, D, 64, 64, levels, _lib.RC_VOL_BF16, fake, None)
        assert r == -1 and b"tensor-core build" in L.rc_last_error()


def test_no_cpu_fallback():
    import raft_stir_b200 as R
    with pytest.raises(RuntimeError, match="CUDA"):
        R.CorrBlock(torch.zeros(1, 8, 16, 16), torch.zeros(1, 8, 16, 16))
    with pytest.raises(RuntimeError, match="CUDA"):
        R.AlternateCorrBlock(torch.zeros(1, 8, 16, 16), torch.zeros(1, 8, 16, 16))
    with pytest.raises(RuntimeError, match="CUDA"):
        R.alt_cuda_corr.forward(torch.zeros(1, 4, 4, 8), torch.zeros(1, 4, 4, 8), torch.zeros(1, 1, 4, 4, 2), 4)


def test_product_does_not_import_oracle():
    pkg = os.path.join(ROOT, "raft_stir_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dirpath, f)).read()
                assert "oracle" not in src, f"{f} mentions the oracle"
    # the GPU-box scripts under tools/ do not use it either (those that do live in tests/perf/)
    for f in os.listdir(os.path.join(ROOT, "tools")):
        if f.endswith((".py", ".sh")):
            src = open(os.path.join(ROOT, "tools", f)).read()
            assert "from oracle" not in src and "import oracle" not in src, f"tools/{f} imports the oracle"


def test_level_view_inverts_the_tile_layout():
    """level_view (what corr_pyramid and the GPU tests read levels through) against
    the element address formula of rc_layout in include/raft_corr.h."""
    import torch
    for vol, shape in ((_lib.RC_VOL_BF16, (2, 13, 21)), (_lib.RC_VOL_F32, (1, 9, 10))):
        B, H, W = shape
        lay = _lib.pyramid_layout(B, H, W, 2, vol)
        buf = torch.arange(int(lay.total_elems), dtype=torch.float32)
        th, tw = lay.tile_h, lay.tile_w
        for l in range(2):
            v = _lib.level_view(buf, lay, l)
            assert tuple(v.shape) == (B * H * W, lay.H[l], lay.W[l])
            for (q, y, x) in ((0, 0, 0), (B * H * W - 1, lay.H[l] - 1, lay.W[l] - 1), (7, 3, 4), (5, 4, 3)):
                if y >= lay.H[l] or x >= lay.W[l]:
                    continue
                want = (lay.offset[l] + q * lay.slab[l] + ((y // th) * lay.tiles_x[l] + x // tw) * th * tw
                        + (y % th) * tw + x % tw)
                assert int(v[q, y, x]) == want


def test_install_into_rebinds_the_reference_names():
    """install_into swaps the names core/raft.py:8 imports and (unless told not to) the two upsampling functions
    (core/raft.py:72-83, core/utils/utils.py:80-82 as imported into core/raft.py)."""
    import types

    import raft_stir_b200 as R

    class FakeRAFT:
        def upsample_flow(self, flow, mask):
            return "reference"

    mod = types.SimpleNamespace(CorrBlock=object, AlternateCorrBlock=object, RAFT=FakeRAFT, upflow8=lambda f: "reference")
    R.install_into(mod, upsampling=False)
    assert mod.CorrBlock is R.CorrBlock and mod.AlternateCorrBlock is R.AlternateCorrBlock
    assert FakeRAFT().upsample_flow(None, None) == "reference" and mod.upflow8(None) == "reference"
    R.install_into(mod)
    assert mod.upflow8 is R.upflow8
    with __import__("pytest").raises(RuntimeError):          # now routed to the library, which rejects CPU tensors
        FakeRAFT().upsample_flow(__import__("torch").zeros(1, 2, 4, 4), __import__("torch").zeros(1, 576, 4, 4))


def test_header_is_plain_c_and_links(tmp_path):
    """include/raft_corr.h compiles as C (no C++, no torch, no CUDA headers needed by a caller) and a C program linked
    against libraftcorr.so can call the host-only entry points."""
    import shutil
    import subprocess

    from raft_stir_b200 import _lib
    gcc = shutil.which("gcc")
    if gcc is None:
        pytest.skip("gcc not available")
    src = tmp_path / "abi.c"
    src.write_text(
        '#include <stdio.h>\n#include "raft_corr.h"\n'
        "int main(void) {\n"
        "  rc_layout lay;\n"
        "  int r = rc_pyramid_layout(2, 55, 128, 4, RC_VOL_BF16, &lay);\n"
        '  printf("%d %d %d %lld %d\\n", rc_abi_version(), r, lay.num_levels, (long long)lay.total_elems, lay.elem_bytes);\n'
        "  return r;\n}\n")
    exe = tmp_path / "abi"
    libdir = os.path.dirname(_lib.LIB_PATH)
    subprocess.run([gcc, "-std=c99", "-Wall", "-Werror", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe),
                    "-L", libdir, "-l:libraftcorr.so", f"-Wl,-rpath,{libdir}"], check=True, capture_output=True)
    out = subprocess.run([str(exe)], check=True, capture_output=True, text=True).stdout.split()
    lay = _lib.pyramid_layout(2, 55, 128, 4, _lib.RC_VOL_BF16)
    assert int(out[0]) == _lib.lib().rc_abi_version() and int(out[1]) == 0
    assert int(out[2]) == 4 and int(out[3]) == int(lay.total_elems) and int(out[4]) == 2
