"""CPU-side checks of the drop-in boundary: the shared library loads, exports every symbol
include/mdb200.h declares, the ctypes prototypes cover exactly that set, struct layouts agree
with the C compiler, and the product path refuses to run without a GPU (no CPU fallback)."""
import ctypes
import os
import re
import subprocess
import sys
import tempfile

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "mdb200.h")


def _declared():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(mdb_[a-z0-9_A-Z]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    from mdb200 import _lib
    names = _declared()
    assert len(names) >= 25
    for n in names:
        assert hasattr(_lib.lib, n), f"libmdb200.so does not export {n}"
    assert sorted(_lib.PROTOTYPES) == names, "ctypes prototypes and header disagree"


def test_no_undeclared_exports():
    from mdb200 import _lib
    out = subprocess.run(["nm", "-D", "--defined-only", _lib.LIB_PATH], capture_output=True, text=True).stdout
    exported = sorted(l.split()[-1] for l in out.splitlines() if " T " in l)
    assert [e for e in exported if not e.startswith("mdb_")] == []
    assert sorted(e for e in exported if e.startswith("mdb_")) == _declared()


def test_struct_layout_matches_c():
    """sizeof/offsetof of the two ABI structs as gcc sees them vs the ctypes mirrors."""
    from mdb200 import _lib
    code = r'''
#include <stdio.h>
#include <stddef.h>
#include "mdb200.h"
int main(void) {
  printf("%zu %zu %zu %zu\n", sizeof(mdb_loss_params), offsetof(mdb_loss_params, alpha),
         offsetof(mdb_loss_params, automasking), offsetof(mdb_loss_params, n_global));
  printf("%zu %zu %zu %zu %zu %zu %zu\n", sizeof(mdb_loss_io), offsetof(mdb_loss_io, rvec),
         offsetof(mdb_loss_io, disparity), offsetof(mdb_loss_io, scale), offsetof(mdb_loss_io, K),
         offsetof(mdb_loss_io, loss), offsetof(mdb_loss_io, vis_warp_loss));
  return 0;
}'''
    with tempfile.TemporaryDirectory() as td:
        c = os.path.join(td, "t.c"); exe = os.path.join(td, "t")
        open(c, "w").write(code)
        subprocess.run(["gcc", "-I", os.path.join(ROOT, "include"), c, "-o", exe], check=True)
        lines = subprocess.run([exe], capture_output=True, text=True, check=True).stdout.split("\n")
    p, io = _lib.LossParams, _lib.LossIO
    assert [int(v) for v in lines[0].split()] == [ctypes.sizeof(p), p.alpha.offset, p.automasking.offset, p.n_global.offset]
    assert [int(v) for v in lines[1].split()] == [ctypes.sizeof(io), io.rvec.offset, io.disparity.offset,
                                                  io.scale.offset, io.K.offset, io.loss.offset, io.vis_warp_loss.offset]


def test_defaults_and_version():
    from mdb200 import _lib
    assert _lib.lib.mdb_version() == 100
    p = _lib.LossParams()
    _lib.lib.mdb_loss_params_default(ctypes.byref(p))
    assert (p.min_depth, p.max_depth, p.disparity_smoothness) == (0.1, 100.0, 1e-3)
    assert abs(p.alpha - 0.85) < 1e-7 and p.automasking == 1 and p.normalize_disparity == 1


def test_workspace_query_and_shape_validation_need_no_gpu():
    from mdb200 import _lib
    io = _lib.LossIO()
    io.W, io.H, io.C, io.L, io.N, io.S, io.Ls = 416, 128, 3, 3, 12, 2, 4
    for k, (w, h) in enumerate([(52, 16), (104, 32), (208, 64), (416, 128)]):
        io.dw[k], io.dh[k] = w, h
    ws = _lib.lib.mdb_train_loss_workspace_bytes(ctypes.byref(io))
    assert ws > 3 * 12 * 416 * 128 * 4  # three low-res scales need a full-res gradient buffer
    io.C = 2
    assert _lib.lib.mdb_train_loss_workspace_bytes(ctypes.byref(io)) == 0
    io.C, io.W = 3, 1
    assert _lib.lib.mdb_train_loss_workspace_bytes(ctypes.byref(io)) == 0
    io.W, io.N = 416, 65536                       # N rides in gridDim.y / z
    assert _lib.lib.mdb_train_loss_workspace_bytes(ctypes.byref(io)) == 0
    io.N, io.W, io.H = 1, 32768, 8192             # W H C L >= 2^31: 32-bit offsets inside an item
    for k in range(4):
        io.dw[k], io.dh[k] = 32768, 8192
    assert _lib.lib.mdb_train_loss_workspace_bytes(ctypes.byref(io)) == 0


@pytest.mark.skipif(__import__("torch").cuda.is_available(), reason="checks the no-GPU behaviour")
def test_product_path_fails_loudly_without_gpu():
    import mdb200
    from mdb200 import _lib
    h = ctypes.c_void_p()
    assert _lib.lib.mdb_create(ctypes.byref(h), 0) == -4  # MDB_E_NO_DEVICE, no CPU fallback
    with pytest.raises(RuntimeError):
        mdb200.handle()


def test_product_code_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "monodepth2.jl_b200")
    for dp, _, fs in os.walk(pkg):
        for f in fs:
            if f.endswith((".py", ".cu", ".cuh", ".inl", ".jl")):
                txt = open(os.path.join(dp, f)).read()
                assert "torch_ref" not in txt and "from oracle" not in txt and "import oracle" not in txt, f


def test_host_loss_default_chunk():
    """HostLoss's pipeline chunk: an eighth of the batch at the bench size, whole small batches in few chunks."""
    import mdb200
    pyr = lambda W, H: [(W // 8, H // 8), (W // 4, H // 4), (W // 2, H // 2), (W, H)]
    assert mdb200.HostLoss.default_chunk(1024, 320, 3, 3, 64, pyr(1024, 320)) == 8
    assert mdb200.HostLoss.default_chunk(1024, 320, 3, 3, 64, pyr(1024, 320), x_format="u8_strip") == 8
    assert mdb200.HostLoss.default_chunk(416, 128, 3, 3, 12, pyr(416, 128)) >= 6
    assert mdb200.HostLoss.default_chunk(8, 8, 1, 3, 2, [(8, 8)]) == 2


def test_bench_reference_arm_runs_on_the_host():
    """`bench.py --impl reference` (the CPU arm the driver runs beside ours) needs no GPU: one JSON line with the
    contract's keys."""
    import json
    import subprocess
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    out = subprocess.run([sys.executable, os.path.join(root, "bench.py"), "--impl", "reference", "--config", "config2",
                          "--steps", "1", "--warmup", "1", "--cpu-sample-items", "2"],
                         capture_output=True, text=True, timeout=600, cwd=root)
    assert out.returncode == 0, out.stderr[-2000:]
    line = json.loads(out.stdout.strip().splitlines()[-1])
    assert line["impl"] == "reference" and line["value"] > 0 and line["unit"] == "warped_px/s"
    assert line["cpu_baseline"]["kind"] == "port" and line["e2e"]["h2d_bytes_per_step"] == 0


# ---- the Julia glue cannot be executed here (no Julia runtime): what CAN be checked statically is checked ----------------
def _strip_c_comments(s):
    return re.sub(r"/\*.*?\*/", "", s, flags=re.S)


def _c_struct_fields(header, name):
    end = re.search(r"\}\s*" + name + r"\s*;", header).start()
    start = header.rfind("{", 0, end)
    body = header[start + 1:end]
    out = []
    for decl in body.split(";"):
        decl = decl.strip()
        if not decl:
            continue
        for part in decl.split(","):                     # "int32_t W, H, C" / "float K[9], invK[9]"
            m = re.search(r"(\w+)\s*(?:\[[^\]]*\])?\s*$", part.strip())
            out.append(m.group(1))
    return out


def _julia_struct_fields(src, name):
    body = re.search(r"^struct " + name + r"\b(.*?)^end", src, flags=re.S | re.M).group(1)
    body = re.sub(r"#.*", "", body)
    return re.findall(r"(\w+)::", body)


def test_julia_struct_mirrors_follow_the_header():
    """LossIO / LossParams / RenderOpts / AdamParams / HostIO in julia/MDB200.jl list the fields of mdb_loss_io / mdb_loss_params /
    mdb_render_opts / mdb_adam_params / mdb_host_io in the header's order (a `ccall` passes them by reference: order IS the ABI)."""
    header = _strip_c_comments(open(os.path.join(ROOT, "include", "mdb200.h")).read())
    jl = open(os.path.join(ROOT, "monodepth2.jl_b200", "julia", "MDB200.jl")).read()
    for c_name, j_name in [("mdb_loss_io", "LossIO"), ("mdb_loss_params", "LossParams"), ("mdb_render_opts", "RenderOpts"),
                           ("mdb_adam_params", "AdamParams"), ("mdb_host_io", "HostIO")]:
        assert _julia_struct_fields(jl, j_name) == _c_struct_fields(header, c_name), (c_name, j_name)


def test_julia_ccalls_name_exported_symbols_with_the_right_arity():
    """Every `ccall((:mdb_xxx, libmdb200), ret, (argtypes...), args...)` in the glue names a function the header declares,
    with as many argument types as the C prototype has parameters."""
    header = _strip_c_comments(open(os.path.join(ROOT, "include", "mdb200.h")).read())
    protos = {}
    for m in re.finditer(r"\b(mdb_\w+)\s*\(([^;{]*?)\)\s*;", header, flags=re.S):
        params = m.group(2).strip()
        protos[m.group(1)] = 0 if params in ("", "void") else len(params.split(","))
    jdir = os.path.join(ROOT, "monodepth2.jl_b200", "julia")
    seen = 0
    for fn in ("MDB200.jl", "MDB200_ops.jl"):
        src = re.sub(r"#.*", "", open(os.path.join(jdir, fn)).read())
        for m in re.finditer(r"ccall\(\(:(\w+),\s*libmdb200\),\s*\w+,\s*\(", src):
            sym, i, depth = m.group(1), m.end(), 1
            j = i
            while depth:                                    # the argument-type tuple, balanced
                depth += {"(": 1, ")": -1}.get(src[j], 0)
                j += 1
            types = src[i:j - 1]
            # top-level commas of the tuple (CuPtr{Cfloat} has none inside; NTuple{9,Cfloat} would: count at brace depth 0)
            n, d = (1 if types.strip().rstrip(",").strip() else 0), 0
            for ch in types.strip().rstrip(","):
                d += {"{": 1, "}": -1, "(": 1, ")": -1}.get(ch, 0)
                n += ch == "," and d == 0
            assert sym in protos, f"{fn}: ccall of undeclared symbol {sym}"
            assert n == protos[sym], f"{fn}: {sym} takes {protos[sym]} parameters, the ccall lists {n} types"
            seen += 1
    assert seen >= 30


def test_strip_geometry_choices_need_no_gpu():
    """mdb_train_loss_strip_geometry is host arithmetic: the launch decomposition the docs quote.  Config 5 (64 x 1024x320, 4
    scales): 37 strips, three tapered segments (205, 92 and 23 rows), 28 416 warps; 8 items (one rank of the 8-GPU run): six
    tapered segments; 12 items at 1024x320 fill exactly one resident wave with whole-height warps: one segment; config 1
    (1 x 416x128, 1 scale): 32 four-row segments so that one wave fills the machine."""
    import mdb200._lib as L

    def geo(W, H, N, Ls):
        io = L.LossIO()
        io.W, io.H, io.C, io.L, io.N, io.S, io.Ls = W, H, 3, 3, N, 2, Ls
        for k in range(Ls):
            io.dw[k], io.dh[k] = max(2, W >> (Ls - 1 - k)), max(2, H >> (Ls - 1 - k))
        out = (ctypes.c_int32 * 4)()
        assert L.lib.mdb_train_loss_strip_geometry(ctypes.byref(io), out) == 0
        return list(out)

    assert geo(1024, 320, 64, 4) == [37, 3, 205, 64 * 4 * 37 * 3]
    assert geo(1024, 320, 8, 4)[:2] == [37, 6]
    assert geo(1024, 320, 12, 4) == [37, 1, 320, 12 * 4 * 37]
    assert geo(416, 128, 1, 1) == [15, 32, 4, 480]
    assert geo(640, 192, 32, 4)[:2] == [23, 3]          # config 3: 1.66 resident waves per segment index
    assert geo(1024, 320, 16, 4)[:2] == [37, 4]         # 1.33 waves: four segments


def test_tapered_segment_boundaries_are_well_formed():
    """The host-side rule that cuts an image into tapered row segments (strip_geom, csrc/photo_fused_host.inl), read back
    through the "strip_seg_dump" test hook: for every forced count k, every exponent and a range of heights the boundaries
    start at 0, end at H, every segment has at least four rows (the kernel's halo is two rows on either side), and the
    segments never grow (long ones are launched first).  Where the image is too short for k segments (H < 16 k) the rule
    falls back to equal segments."""
    import mdb200._lib as L
    dump = (ctypes.c_int32 * 10)()
    L.lib.mdb_debug_option(b"strip_seg_dump", ctypes.addressof(dump))
    try:
        for H in (32, 37, 64, 100, 128, 192, 320, 375, 1080):
            for k in range(2, 9):
                for pow10 in (5, 10, 20, 40):
                    L.lib.mdb_debug_option(b"strip_taper", k)
                    L.lib.mdb_debug_option(b"strip_taper_pow10", pow10)
                    io = L.LossIO()
                    io.W, io.H, io.C, io.L, io.N, io.S, io.Ls = 416, H, 3, 3, 5, 2, 1
                    io.dw[0], io.dh[0] = 416, H
                    out = (ctypes.c_int32 * 4)()
                    assert L.lib.mdb_train_loss_strip_geometry(ctypes.byref(io), out) == 0
                    if H < 16 * k:
                        assert dump[0] == 0, (H, k)
                        continue
                    assert dump[0] == 1 and out[1] == k, (H, k, pow10)
                    y = [dump[1 + i] for i in range(k + 1)]
                    assert y[0] == 0 and y[k] == H, (H, k, pow10, y)
                    rows = [b - a for a, b in zip(y, y[1:])]
                    assert min(rows) >= 4, (H, k, pow10, y)
                    assert all(a >= b for a, b in zip(rows, rows[1:])), (H, k, pow10, y)
                    assert out[2] == rows[0] and out[3] == 5 * 15 * k
        # the shipped exponent at config 5: 205 / 92 / 23 rows
        L.lib.mdb_debug_option(b"strip_taper", 0)
        L.lib.mdb_debug_option(b"strip_taper_pow10", 0)
        io = L.LossIO()
        io.W, io.H, io.C, io.L, io.N, io.S, io.Ls = 1024, 320, 3, 3, 64, 2, 4
        for s in range(4):
            io.dw[s], io.dh[s] = 1024 >> (3 - s), 320 >> (3 - s)
        out = (ctypes.c_int32 * 4)()
        assert L.lib.mdb_train_loss_strip_geometry(ctypes.byref(io), out) == 0
        assert [dump[i] for i in range(5)] == [1, 0, 205, 297, 320]
    finally:
        L.lib.mdb_debug_option(b"strip_taper", 0)
        L.lib.mdb_debug_option(b"strip_taper_pow10", 0)
        L.lib.mdb_debug_option(b"strip_seg_dump", 0)
