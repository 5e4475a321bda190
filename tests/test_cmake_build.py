"""The CMake hook of the drop-in (north star: "the build gains an sm_100a CUDA target next to cmake_modules/FindG2O.cmake"; the
reference's own build is CMakeLists.txt:28-30 + cmake_modules/FindG2O.cmake:1-113).  CPU: configure and build the repo's CMake project
(nvcc cross-compiles sm_100a without a GPU), check the product library it produced, then let a small consumer project find it through
cmake_modules/FindBAEngineCUDA.cmake and link a C program against the C-ABI."""
import ctypes as C
import os
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
NVCC = "/usr/local/cuda/bin/nvcc"


@pytest.fixture(scope="module")
def cmake_build(tmp_path_factory):
    if not shutil.which("cmake") or not shutil.which("ninja") or not os.path.exists(NVCC):
        pytest.skip("cmake / ninja / nvcc not available")
    build = tmp_path_factory.mktemp("cmake_build")
    subprocess.check_call(["cmake", "-S", ROOT, "-B", str(build), "-G", "Ninja", f"-DCMAKE_CUDA_COMPILER={NVCC}", "-DBA_BUILD_ORACLE=OFF"],
                          stdout=subprocess.DEVNULL)
    subprocess.check_call(["cmake", "--build", str(build), "--target", "ba_engine", "ba_synth"], stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    return build


def test_cmake_builds_the_sm100a_engine(cmake_build):
    from test_capi_cpu import declared_symbols
    lib_path = os.path.join(str(cmake_build), "libba_engine.so")
    assert os.path.exists(lib_path)
    lib = C.CDLL(lib_path)
    for n in declared_symbols("ba_engine.h"):
        assert hasattr(lib, n), n
    lib.ba_backend.restype = C.c_char_p
    assert lib.ba_backend() == b"cuda-sm100a"
    # sm_100a code only: no other architecture, no PTX fallback
    elf = subprocess.check_output(["/usr/local/cuda/bin/cuobjdump", "-lelf", lib_path], text=True)
    archs = {line.split(".")[-2] for line in elf.split() if line.endswith(".cubin")}
    assert archs == {"sm_100a"}, archs
    ptx = subprocess.run(["/usr/local/cuda/bin/cuobjdump", "-lptx", lib_path], capture_output=True, text=True).stdout
    assert ".ptx" not in ptx


def test_find_module_links_a_consumer(cmake_build, tmp_path):
    """What the reference's CMakeLists.txt would do after dropping find_package(G2O): find_package(BAEngineCUDA) + link."""
    src = tmp_path / "consumer"
    src.mkdir()
    (src / "main.c").write_text('#include <stdio.h>\n#include "ba_engine.h"\nint main(void){ba_options o; if (ba_options_default(&o)) return 1;'
                                'printf("%s %d %g\\n", ba_backend(), o.max_num_iterations, o.robust_a); return 0;}\n')
    (src / "CMakeLists.txt").write_text(
        "cmake_minimum_required(VERSION 3.24)\nproject(consumer LANGUAGES C)\n"
        f"list(APPEND CMAKE_MODULE_PATH {ROOT}/cmake_modules)\n"
        "find_package(BAEngineCUDA REQUIRED)\nadd_executable(consumer main.c)\ntarget_link_libraries(consumer PRIVATE BAEngineCUDA::ba_engine)\n")
    # BAENGINE_ROOT = the source tree (headers) ; the library comes from the build tree through CMAKE_LIBRARY_PATH
    out = tmp_path / "consumer_build"
    subprocess.check_call(["cmake", "-S", str(src), "-B", str(out), "-G", "Ninja", f"-DBAENGINE_ROOT={ROOT}",
                           f"-DBAEngineCUDA_LIBRARY={cmake_build}/libba_engine.so"], stdout=subprocess.DEVNULL)
    subprocess.check_call(["cmake", "--build", str(out)], stdout=subprocess.DEVNULL)
    env = dict(os.environ, LD_LIBRARY_PATH=f"{cmake_build}:/usr/local/cuda/lib64:" + os.environ.get("LD_LIBRARY_PATH", ""))
    got = subprocess.check_output([str(out / "consumer")], env=env, text=True).split()
    assert got == ["cuda-sm100a", "50", "4"]
