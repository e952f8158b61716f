"""CPU tests of the drop-in boundary: the C-ABI library loads and exports every symbol include/vr5.h declares, the POD
mirrors have the reference's layout, the C++ adaptor compiles against the reference's REAL headers, and without a GPU
the product fails loudly (no CPU fallback)."""
import ctypes as C
import os
import re
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = "/root/reference/VolumeRendering"
CSRC = os.path.join(ROOT, "volume-rendering_b200", "csrc")


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "vr5.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(vr5_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol(R):
    import vr5_loader
    pkg = vr5_loader.load()
    lib = C.CDLL(pkg.LIB_PATH)
    names = declared_symbols()
    assert len(names) >= 35
    missing = [n for n in names if not hasattr(lib, n)]
    assert not missing, missing
    # and the ctypes binding table covers the same set
    assert pkg.lib().vr5_version() == 1
    assert b"GPURenderer5" in pkg.lib().vr5_name()


def test_no_torch_or_cxx_types_in_the_abi():
    text = re.sub(r"/\*.*?\*/", "", open(os.path.join(ROOT, "include", "vr5.h")).read(), flags=re.S)  # declarations only
    assert "torch" not in text and "std::" not in text and "cudaStream_t" not in text


def test_pod_layout_matches_python_mirrors(R):
    import oracle
    assert C.sizeof(R.View) == 68 and C.sizeof(oracle.Raycaster) == 160 and C.sizeof(oracle.Model) == 32
    assert oracle.Raycaster.view.offset == 32 and oracle.Raycaster.transfer_fn.offset == 104
    assert oracle.Raycaster.ray_step.offset == 112 and oracle.Raycaster.esl_volume.offset == 128
    assert oracle.Raycaster.esl_block_dims.offset == 136 and oracle.Raycaster.esl_block_size.offset == 140
    assert oracle.Raycaster.light_kd.offset == 152
    assert C.sizeof(R.Params) == 112


def test_pod_header_compiles_as_c_and_cxx(tmp_path):
    src = tmp_path / "t.c"
    src.write_text('#include "vr5.h"\nint main(void){ vr5_params p; (void)p; return (int)sizeof(vr5_raycaster) - 160; }\n')
    for cc, std in (("gcc", "-std=c11"), ("g++", "-std=c++17")):
        exe = tmp_path / ("t_" + cc)
        subprocess.run([cc, std, "-x", "c" if cc == "gcc" else "c++", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe)], check=True)
        assert subprocess.run([str(exe)]).returncode == 0


@pytest.mark.skipif(not os.path.isdir(REF), reason="needs the reference headers")
def test_adaptor_compiles_against_reference_headers(tmp_path):
    """GPURenderer5.cpp built the way a VolR maintainer would: against Renderer.h / RaycasterBase.h / ViewBase.h /
    ModelBase.h of the reference itself, plus layout asserts tying vr5_pod.h to those real structs."""
    check = tmp_path / "layout_check.cpp"
    check.write_text(
        '#include <stddef.h>\n#include "Renderer.h"\n#include "vr5_pod.h"\n'
        "static_assert(sizeof(Model)==sizeof(vr5_model) && sizeof(View)==sizeof(vr5_view) && sizeof(Raycaster)==sizeof(vr5_raycaster), \"sizes\");\n"
        "static_assert(offsetof(Model,size)==offsetof(vr5_model,size) && offsetof(Model,dims)==offsetof(vr5_model,dims) && offsetof(Model,min_bound)==offsetof(vr5_model,min_bound), \"Model\");\n"
        "static_assert(offsetof(View,origin)==offsetof(vr5_view,origin) && offsetof(View,direction)==offsetof(vr5_view,direction) && offsetof(View,right_plane)==offsetof(vr5_view,right_plane) && offsetof(View,up_plane)==offsetof(vr5_view,up_plane) && offsetof(View,light_pos)==offsetof(vr5_view,light_pos) && offsetof(View,perspective)==offsetof(vr5_view,perspective), \"View\");\n"
        "static_assert(offsetof(Raycaster,view)==offsetof(vr5_raycaster,view) && offsetof(Raycaster,transfer_fn)==offsetof(vr5_raycaster,transfer_fn) && offsetof(Raycaster,ray_step)==offsetof(vr5_raycaster,ray_step) && offsetof(Raycaster,ray_threshold)==offsetof(vr5_raycaster,ray_threshold) && offsetof(Raycaster,esl)==offsetof(vr5_raycaster,esl) && offsetof(Raycaster,esl_volume)==offsetof(vr5_raycaster,esl_volume) && offsetof(Raycaster,esl_block_dims)==offsetof(vr5_raycaster,esl_block_dims) && offsetof(Raycaster,esl_block_size)==offsetof(vr5_raycaster,esl_block_size) && offsetof(Raycaster,light_kd)==offsetof(vr5_raycaster,light_kd), \"Raycaster\");\n"
        "static_assert(TF_SIZE==VR5_TF_SIZE && TF_RATIO==VR5_TF_RATIO && ESL_VOLUME_SIZE==VR5_ESL_VOLUME_SIZE && ESL_VOLUME_DIMS==VR5_ESL_VOLUME_DIMS && ESL_MIN_BLOCK_SIZE==VR5_ESL_MIN_BLOCK_SIZE, \"constants\");\n"
        "int main(){return 0;}\n")
    inc = ["-I", REF, "-I", "/usr/local/cuda/include", "-I", os.path.join(ROOT, "include")]
    subprocess.run(["g++", "-std=c++17", "-w", "-c", str(check), "-o", str(tmp_path / "c.o")] + inc, check=True)
    subprocess.run(["g++", "-std=c++17", "-w", "-DVR5_WITH_VOLR_HEADERS", "-c", os.path.join(CSRC, "GPURenderer5.cpp"),
                    "-o", str(tmp_path / "g5.o")] + inc, check=True)


def test_fails_loudly_without_a_gpu(R):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(Exception) as e:
        R.GPURenderer5(0)
    assert "no CPU fallback" in str(e.value)


def test_product_never_imports_the_oracle():
    """The product path may not import, link or execute anything under oracle/ (a CPU fallback would void parity)."""
    pkg_dir = os.path.join(ROOT, "volume-rendering_b200")
    for dirpath, _, files in os.walk(pkg_dir):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h")) or f == "Makefile":
                text = open(os.path.join(dirpath, f), errors="ignore").read()
                assert "import oracle" not in text and "from oracle" not in text and "liboracle" not in text, f
                assert "oracle_port" not in text and "libvolr_ref" not in text, f
    out = subprocess.run(["ldd", os.path.join(pkg_dir, "libvr5.so")], capture_output=True, text=True).stdout
    assert "oracle" not in out and "volr_ref" not in out


def test_kernel_has_no_fused_multiply_add_in_ray_math():
    """-fmad=false is part of the parity contract; the only FFMA in the parity kernels may come from the IEEE division /
    square-root sequences (FCHK/MUFU neighbourhood), never from packed FFMA2 contraction. The one place a packed FMA is
    legitimate is the contracted trilinear policy (tri_kernel<..., FMA=true>), whose fusions are explicit."""
    lib = os.path.join(ROOT, "volume-rendering_b200", "libvr5.so")
    sass = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
    assert "cast_kernel" in sass
    with_ffma2 = set()
    fn = None
    for line in sass.splitlines():
        if "Function :" in line:
            fn = line.split("Function :")[1].strip()
        elif "FFMA2" in line and fn:
            with_ffma2.add(fn)
    # mangled template arguments of tri_kernel: COUNT, FAST, SHADE, FMA, MINB, DEPTH, ZS -> ...ILb?ELb?ELb?ELb1ELi?ELi?ELb?EE...
    bad = [f for f in with_ffma2 if not re.search(r"tri_kernelILb[01]ELb[01]ELb[01]ELb1ELi\d+ELi\d+ELb[01]EE", f)]
    assert not bad, bad
    assert any("tri_kernel" in f for f in with_ffma2)  # the contracted march does use the packed FMA
    mk = open(os.path.join(CSRC, "Makefile")).read()
    assert "-fmad=false" in mk and "use_fast_math" not in mk.replace("Never --use_fast_math", "")


def test_predicted_empty_space_leap_acceptance_rule(tmp_path):
    """vr5_cast.cuh predicts the leap count floor(dk / step) of RaycasterBase.h:67-85 with reciprocal multiplications and accepts the
    prediction only at a stated distance from the neighbouring integers; tools/check_esl_predict.c replays the rule on the host
    against the reference's divisions, random and adversarial (quotients within a few ulps of an integer) operands."""
    exe = tmp_path / "check_esl"
    subprocess.run(["gcc", "-O2", "-ffp-contract=off", os.path.join(ROOT, "tools", "check_esl_predict.c"), "-o", str(exe), "-lm"], check=True)
    out = subprocess.run([str(exe), "5"], check=True, capture_output=True, text=True).stdout
    assert " 0 mismatches" in out, out


def test_constant_increment_of_the_ray_parameter(tmp_path):
    """The leaps of the zero-skip kernels (vr5_cast.cuh, try_leap) replace n sequential k += step of CPURenderer.cpp:37 by
    bits(k) + n * M inside one binade; tools/check_kleap.c walks random and adversarial (k, step) pairs step by step and compares."""
    exe = tmp_path / "check_kleap"
    subprocess.run(["gcc", "-O2", "-ffp-contract=off", os.path.join(ROOT, "tools", "check_kleap.c"), "-o", str(exe), "-lm"], check=True)
    out = subprocess.run([str(exe), "20000"], check=True, capture_output=True, text=True).stdout
    assert "mismatches 0" in out and "steps compared 0" not in out, out


def test_leap_geometry_against_the_reference_arithmetic(tmp_path):
    """tools/check_leap_geometry.c replays the leap rule of the zero-skip kernels (brick look-up, shrunk exit times, brick-to-brick walk
    over clean crossings, floor(t / inc), binade cap) on random and adversarial rays and checks that every sample a leap leaves out
    reads, in the reference's own fp32 arithmetic (ModelBase.h:17-23 on origin + direction * k), a voxel of a brick the walk accepted,
    and that the ray parameter after the leap is the reference's. Without the margin the same check must FIND violations."""
    exe = tmp_path / "check_leap_geometry"
    subprocess.run(["gcc", "-O1", "-ffp-contract=off", os.path.join(ROOT, "tools", "check_leap_geometry.c"), "-o", str(exe), "-lm"], check=True)
    out = subprocess.run([str(exe), "150000"], check=True, capture_output=True, text=True).stdout
    assert "violations 0 " in out and "leaps 0," not in out, out
    bad = subprocess.run([str(exe), "150000"], capture_output=True, text=True, env=dict(os.environ, CLG_MARGIN_SCALE="0"))
    assert bad.returncode != 0 and "VOXEL VIOLATION" in bad.stdout, bad.stdout[-500:]
