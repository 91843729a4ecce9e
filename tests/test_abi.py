"""CPU tests of the drop-in boundary: the C-ABI library loads, exports every symbol
include/tlfea_b200.h declares, the ctypes table is in sync with the header, and the product
refuses to run without a CUDA device (no CPU fallback)."""
import ctypes as C
import os
import re

import pytest

from conftest import ROOT, _has_gpu

HEADER = os.path.join(ROOT, "include", "tlfea_b200.h")


def _declared():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(tlfea_[A-Za-z0-9_]+)\s*\(", src)))


def test_header_symbols_exported_and_bound():
    import tlfea
    from tlfea import _lib

    names = _declared()
    assert len(names) > 60
    raw = C.CDLL(_lib.LIB_PATH)
    for n in names:
        assert hasattr(raw, n), f"{n} declared in the header but not exported"
    assert sorted(_lib.exported_symbols()) == names
    assert tlfea.lib.tlfea_version() >= 100


def test_params_struct_layout():
    from tlfea import NewtonParams

    # SyncedNewtonParams {double x4; int x2; double} -> 48 bytes (SyncedNewton.cuh:29-33)
    assert C.sizeof(NewtonParams) == 48
    assert NewtonParams.time_step.offset == 40


@pytest.mark.skipif(_has_gpu(), reason="checks the no-device error path")
def test_no_cpu_fallback():
    import tlfea

    with pytest.raises(tlfea.TlfeaError) as ei:
        tlfea.GPU_FEAT10_Data(4, 20)
    assert ei.value.code == -4  # TLFEA_ERR_NO_DEVICE


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "total-lagrangian-fea_b200")
    for dp, _, fns in os.walk(pkg):
        for fn in fns:
            if fn.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
                txt = open(os.path.join(dp, fn)).read()
                assert "oracle" not in txt.lower() or fn == "README.md", f"{fn} mentions the oracle"


def test_cpp_shim_compiles_and_links_against_the_abi(tmp_path):
    """include/tlfea_shim.hpp (reference class names forwarding to the C ABI) builds with a plain host compiler and
    every symbol it uses resolves in libtlfea_b200.so (no run: there is no GPU here)."""
    import os
    import shutil
    import subprocess

    gxx = shutil.which("g++")
    if gxx is None:
        pytest.skip("no g++")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    lib_dir = os.path.join(root, "total-lagrangian-fea_b200", "lib")
    src = tmp_path / "driver.cc"
    src.write_text(
        '#include "tlfea_shim.hpp"\n'
        "int main(int argc, char **) {\n"
        "  if (argc < 99) return 0;  // link check only\n"
        "  GPU_FEAT10_Data t(1, 10); GPU_ANCF3243_Data b(3, 2); GPU_ANCF3443_Data p(4, 1);\n"
        "  SyncedNewtonSolver s(&t, 0); SyncedNewtonParams prm{1e-4, 0, 1e-6, 1e14, 5, 10, 1e-3};\n"
        "  s.SetParameters(&prm); s.Setup(); s.AnalyzeHessianSparsity(); s.Solve();\n"
        "  std::vector<double> f; t.RetrieveInternalForceToCPU(f); b.CalcDsDuPre(); p.CalcMassMatrix();\n"
        "  std::vector<int> o, c; std::vector<double> v; b.RetrieveMassCSRToCPU(o, c, v); b.UpdateNodalFixed(o);\n"
        "  return (int)f.size();\n}\n")
    exe = tmp_path / "driver"
    subprocess.check_call([gxx, "-std=c++17", "-I", os.path.join(root, "include"), str(src), "-o", str(exe),
                           "-L", lib_dir, "-ltlfea_b200", "-Wl,-rpath," + lib_dir, "-Wl,--unresolved-symbols=ignore-in-shared-libs"])
    assert subprocess.call([str(exe)], env=dict(os.environ, LD_LIBRARY_PATH=lib_dir)) == 0
    # the beam_sag example driver builds against the same header and library
    ex = tmp_path / "beam_sag"
    subprocess.check_call([gxx, "-std=c++17", "-I", os.path.join(root, "include"), os.path.join(root, "examples", "beam_sag.cc"),
                           "-o", str(ex), "-L", lib_dir, "-ltlfea_b200", "-Wl,-rpath," + lib_dir,
                           "-Wl,--unresolved-symbols=ignore-in-shared-libs"])
    # the bunny driver (Eigen-style containers through the shim's templates, examples/mini_matrix.hpp) builds too
    bn = tmp_path / "bunny"
    subprocess.check_call([gxx, "-std=c++17", "-Wall", "-Werror", "-I", os.path.join(root, "include"),
                           "-I", os.path.join(root, "examples"), os.path.join(root, "examples", "bunny_newton.cc"),
                           "-o", str(bn), "-L", lib_dir, "-ltlfea_b200", "-Wl,-rpath," + lib_dir,
                           "-Wl,--unresolved-symbols=ignore-in-shared-libs"])


def test_host_thread_setting_is_accepted_without_a_gpu():
    """tlfea_set_host_threads only touches the OpenMP team of the host-side analyses: callable on the CPU box."""
    from tlfea._lib import lib

    assert lib.tlfea_set_host_threads(2) == 0
    assert lib.tlfea_set_host_threads(0) == 0
    assert lib.tlfea_set_host_threads(-1) != 0 and b"thread count" in lib.tlfea_last_error()
