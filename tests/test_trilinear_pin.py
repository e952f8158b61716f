"""External parity anchor of the trilinear sampler policies.

Reference = the reference's OWN GPURenderer4.cu, compiled unmodified for sm_100a (oracle/ref_gpu4_texobj.cu maps its legacy
texture references onto texture objects): tex3D on the uint8 volume with cudaFilterModeLinear / normalised coordinates /
cudaReadModeNormalizedFloat / clamp, tex1D on the 128-entry float4 transfer function with linear filtering, nvcc's default
FMA contraction - executed by the B200's texture unit. VR5_SAMPLER_TRILINEAR and VR5_SAMPLER_TRILINEAR_FMA evaluate the
same formulas in software fp32. The texture unit filters with 8 fractional weight bits (CUDA C Programming Guide, "Linear
Filtering": the interpolation weights are 9-bit fixed point), the software policies with 24-bit mantissas, so the frames
are NOT bit-equal; the tolerance below is what tools/ref4_pin.py measured on this hardware (profiles/r02_ref4_pin.json holds
the histograms), stated per class of scene:

  * no threshold decision in play (semi-transparent TF: no early termination, no shading):
        <= 1 LSB on >= TOL_SMOOTH_WITHIN1 of the pixels, <= TOL_SMOOTH_MAX LSB max
  * decisions in play (default / opaque TF: a sample whose alpha sits near the 0.05 shading guard, a ray whose opacity sits
    near the 0.95 threshold, or a leading empty-space test can fall the other way when the filtered value moves by 1/256 of
    a voxel difference - that changes a whole sample's contribution):
        <= 2 LSB on >= TOL_EDGY_WITHIN2 and <= 4 LSB on >= TOL_EDGY_WITHIN4 of the pixels; the maximum is reported, not bounded
        (one flipped first-hit sample of an opaque TF changes a pixel completely; measured: 4..28 LSB with the default TF, 111
        with the opaque one)

and, as the discriminating control, the nearest-neighbour parity policy is much farther from GPURenderer4 than either
trilinear policy on the same frames."""
import pytest

import oracle
from common import golden_scene

TOL_SMOOTH_WITHIN1, TOL_SMOOTH_MAX = 0.999, 2
TOL_EDGY_WITHIN2, TOL_EDGY_WITHIN4 = 0.97, 0.98   # measured minima: 0.9807 / 0.9827 (synthB64_opaque_ortho)

SMOOTH = ["synthA64_semi_ortho", "synthA64_semi_persp"]
EDGY = ["bucky_ortho_0_0_0", "bucky_persp_-45_-45_0", "bucky_ortho_90_0_0", "synthA64_default_ortho", "synthA64_default_persp",
        "synthA96_default_step1.7", "synthB64_default_noesl", "synthB64_default_thr1_kd0", "synthB64_opaque_ortho", "synthB64_opaque_persp"]


def _compare(name, golden, port, volumes, R):
    import sys
    import os
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tools"))
    import ref4_pin
    if not oracle.RefGpu.available():
        pytest.skip("oracle/_ref/libvolr_ref_gpu.so not built (needs /root/reference at build time)")
    case = next(c for c in golden["cases"] if c["name"] == name)
    sc = golden_scene(port, case, volumes(case["volume"]))
    return ref4_pin.compare_case(R, oracle, sc)


@pytest.mark.gpu
@pytest.mark.parametrize("name", SMOOTH)
def test_trilinear_policies_match_gpurenderer4_smooth(name, golden, port, volumes, R, record_property):
    rep = _compare(name, golden, port, volumes, R)
    record_property("lsb_histograms", rep)
    for sampler in ("trilinear", "trilinear-fma"):
        r = rep[sampler]
        assert r["within1"] >= TOL_SMOOTH_WITHIN1 and r["max_lsb"] <= TOL_SMOOTH_MAX, (name, sampler, r)
    assert rep["hybridw"]["within1"] < rep["trilinear-fma"]["within1"]   # the nearest policy is not GPURenderer4's sampler


@pytest.mark.gpu
@pytest.mark.parametrize("name", EDGY)
def test_trilinear_policies_match_gpurenderer4_with_decisions(name, golden, port, volumes, R, record_property):
    rep = _compare(name, golden, port, volumes, R)
    record_property("lsb_histograms", rep)
    for sampler in ("trilinear", "trilinear-fma"):
        r = rep[sampler]
        assert r["within2"] >= TOL_EDGY_WITHIN2 and r["within4"] >= TOL_EDGY_WITHIN4, (name, sampler, r)
        assert sum(r["mean_abs_per_channel"]) < sum(rep["hybridw"]["mean_abs_per_channel"]), (name, sampler)


@pytest.mark.gpu
def test_trilinear_policies_match_gpurenderer4_256_cubed_orbit(R, record_property):
    """BASELINE configs[1] at a smaller viewport: synthA 256^3, semi-transparent TF, three orbit poses (none parallel to a cube
    face: there the FMA-contracted ray / box clip of the reference's GPU renderers drops a pixel column that every sampler of
    this library - following CPURenderer's uncontracted clip - renders; profiles/r02_ref4_pin.md, c2_orbit). Smooth-scene
    tolerance: <= 1 LSB on >= 99.9 % of the pixels, <= 2 LSB max."""
    import os
    import sys
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tools"))
    import bench
    import ref4_pin
    if not oracle.RefGpu.available():
        pytest.skip("oracle/_ref/libvolr_ref_gpu.so not built (needs /root/reference at build time)")
    wl = dict(bench.WORKLOADS["c2"], w=640, h=360)
    cpu = bench.CpuReference(wl)
    scs = [cpu.scene((20.0, ang, 0.0)) for ang in (37.0, 111.0, 200.0)]
    rep = ref4_pin.compare_case(R, oracle, scs[0], rcs=[s.rc for s in scs])
    record_property("lsb_histograms", rep)
    for sampler in ("trilinear", "trilinear-fma"):
        r = rep[sampler]
        assert r["lit_pixels"] > 100_000 and r["within1"] >= TOL_SMOOTH_WITHIN1 and r["max_lsb"] <= TOL_SMOOTH_MAX, (sampler, r)
    assert rep["hybridw"]["within1"] < 0.99   # the nearest policy is measurably not GPURenderer4's sampler
