"""Build recipe for oracle/_ref: the UNMODIFIED reference graph-builder kernel.

TEST INFRASTRUCTURE ONLY.  This compiles the reference's own CUDA extension
(`/root/reference/src/dagr/graph/ev_graph.cu` + `spiral.h`, pybind module name
`ev_graph_cuda`, as `setup.py:9-12` of the reference names it) for sm_100a, from the
sources where they lie.  No reference source is copied into this repository; the only
output is `oracle/_ref/ev_graph_cuda*.so` (git-ignored, but it travels to the GPU box).

The resulting module is a *second oracle* for `edge_index`: on a B200 the GPU tests run
`fill_edges_cuda` / `insert_in_queue_cuda` on the same inputs as our kernels and as the C
restatement in `oracle/evg_oracle.c`, which pins the restatement to the real reference.

Usage:  python oracle/build_ref.py        (needs /root/reference; ~5 min because of torch/extension.h)
"""
import os
import subprocess
import sys
import sysconfig

REF_SRC = "/root/reference/src/dagr/graph/ev_graph.cu"
REF_ASY_SRC = "/root/reference/src/dagr/asynchronous/asy_tools/main.cu"     # pybind module `asy_tools` (setup.py:9-12)
HERE = os.path.dirname(os.path.abspath(__file__))
OUT_DIR = os.path.join(HERE, "_ref")
OUT = os.path.join(OUT_DIR, "ev_graph_cuda.so")
OUT_ASY = os.path.join(OUT_DIR, "asy_tools.so")


def build_asy(force: bool = False) -> str:
    """The reference's asy_tools extension (masked_lin, masked_lin_no_bias, masked_isdiff, masked_inplace_BN), unmodified."""
    return build(force, REF_ASY_SRC, OUT_ASY, "asy_tools")


def build(force: bool = False, REF_SRC: str = REF_SRC, OUT: str = OUT, name: str = "ev_graph_cuda") -> str:
    if not os.path.exists(REF_SRC):
        if os.path.exists(OUT):
            return OUT
        raise FileNotFoundError(f"{REF_SRC} not present (the reference is only mounted in the build container)")
    if os.path.exists(OUT) and not force and os.path.getmtime(OUT) > os.path.getmtime(REF_SRC):
        return OUT
    import torch
    from torch.utils import cpp_extension

    os.makedirs(OUT_DIR, exist_ok=True)
    inc = cpp_extension.include_paths("cuda") if hasattr(cpp_extension, "include_paths") else []
    inc = list(inc) + [sysconfig.get_paths()["include"], os.path.dirname(REF_SRC)]
    torch_lib = os.path.join(os.path.dirname(torch.__file__), "lib")
    cmd = ["nvcc", "-shared", "-Xcompiler", "-fPIC", "-O2", "-std=c++17",
           "-gencode", "arch=compute_100a,code=sm_100a",
           f"-DTORCH_EXTENSION_NAME={name}", "-DTORCH_API_INCLUDE_EXTENSION_H",
           f"-D_GLIBCXX_USE_CXX11_ABI={int(torch._C._GLIBCXX_USE_CXX11_ABI)}",
           "-w"]
    for i in inc:
        cmd += ["-I", i]
    cmd += ["-o", OUT, REF_SRC, "-L", torch_lib,
            "-lc10", "-lc10_cuda", "-ltorch_cpu", "-ltorch_cuda", "-ltorch", "-ltorch_python",
            "-Xlinker", f"-rpath={torch_lib}"]
    print(" ".join(cmd), flush=True)
    subprocess.check_call(cmd)
    return OUT


if __name__ == "__main__":
    print(build(force="--force" in sys.argv))
    print(build_asy(force="--force" in sys.argv))
