"""Host logic of the device structure builders, checked without a GPU: the per-channel routine the Subgrid kernels run
(populate_cell, csrc/ptb_grid_gpu.h) is host-compilable on purpose — the SAT code is one source for both sides
(csrc/ptb_sat.h) — so g++ builds a harness that fills every channel with it and compares the packed pools with the host
builder's, byte for byte."""
import os
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CSRC = os.path.join(ROOT, "bachelor-thesis-pathtracer_b200", "csrc")


def test_subgrid_channel_routine_reproduces_the_host_builder(tmp_path):
    exe = str(tmp_path / "subgrid_cells_check")
    cmd = ["g++", "-std=c++17", "-O2", "-ffp-contract=off", "-pthread", "-I", CSRC, "-I", os.path.join(ROOT, "include"),
           os.path.join(ROOT, "tests", "native", "subgrid_cells_check.cpp"), os.path.join(CSRC, "ptb_grid_build.cpp"),
           os.path.join(CSRC, "ptb_bvh_build.cpp"), "-o", exe]
    subprocess.run(cmd, check=True, stdout=subprocess.PIPE, stderr=subprocess.PIPE)
    out = subprocess.run([exe], stdout=subprocess.PIPE, text=True, timeout=600)
    lines = [l for l in out.stdout.splitlines() if l.startswith("n=")]
    assert out.returncode == 0 and len(lines) == 5 and all(l.endswith("IDENTICAL") for l in lines), out.stdout


def _build_and_run(tmp_path, name):
    exe = str(tmp_path / name)
    cmd = ["g++", "-std=c++17", "-O2", "-ffp-contract=off", "-pthread", "-I", CSRC, "-I", os.path.join(ROOT, "include"),
           os.path.join(ROOT, "tests", "native", name + ".cpp"), os.path.join(CSRC, "ptb_grid_build.cpp"),
           os.path.join(CSRC, "ptb_bvh_build.cpp"), "-o", exe]
    subprocess.run(cmd, check=True, stdout=subprocess.PIPE, stderr=subprocess.PIPE)
    return subprocess.run([exe], stdout=subprocess.PIPE, text=True, timeout=600)


def test_level_synchronous_subbeam_builder_reproduces_the_host_builder(tmp_path):
    """csrc/ptb_beam_levels.h — the functors ptb_beam_gpu.cu launches as kernels plus the driver around them — run with
    plain loops: beam records, leaf slices and their numbering equal build_subbeams' depth-first output byte for byte."""
    out = _build_and_run(tmp_path, "subbeam_levels_check")
    lines = [l for l in out.stdout.splitlines() if l.startswith("n=")]
    assert out.returncode == 0 and len(lines) == 7 and all(l.endswith("IDENTICAL") for l in lines), out.stdout
