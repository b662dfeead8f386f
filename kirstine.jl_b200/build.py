"""Builds libkirstine_b200.so in-tree with nvcc for sm_100a (cross-compiles without a GPU).

Usage: python build.py [--force] [--verbose]
Called by __graft_entry__.build(); the resulting .so is git-ignored but travels to the GPU box.
"""
import concurrent.futures as cf
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OBJ = os.path.join(HERE, "build")
LIB = os.path.join(HERE, "libkirstine_b200.so")
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
COMMON = ["-O3", "-std=c++17", "-lineinfo", "-Xcompiler", "-fPIC", "-Xcompiler", "-fvisibility=hidden"]
# per-file extras: the swarm move must round exactly like the scalar reference arithmetic
EXTRA = {"kst_pso.cu": ["-fmad=false"]}
SOURCES = ["kst_api.cu", "kst_comm.cu", "kst_objective.cu", "kst_obj_inst_a.cu", "kst_obj_inst_b.cu", "kst_obj_inst_c.cu", "kst_obj_inst_d.cu",
           "kst_obj_inst_e.cu", "kst_obj_inst_f.cu", "kst_gateaux.cu", "kst_pso.cu", "kst_exchange.cu"]


def _newer(target, deps):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps)


def _headers():
    hs = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cuh", ".h", ".inc"))]
    hs.append(os.path.join(os.path.dirname(HERE), "include", "kirstine_b200.h"))
    hs.append(os.path.abspath(__file__))
    return hs


def _compile(src, verbose):
    obj = os.path.join(OBJ, src.replace(".cu", ".o"))
    cmd = [NVCC, *ARCH, *COMMON, *EXTRA.get(src, []), "-c", os.path.join(CSRC, src), "-o", obj]
    if verbose:
        cmd.insert(1, "-Xptxas=-v")
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("nvcc failed for %s:\n%s\n%s" % (src, r.stdout, r.stderr))
    return src, r.stderr


def build_variant(name, defines):
    """Tuning helper: build lib<name>.so with extra -D flags into its own object directory."""
    global OBJ, LIB, COMMON
    saved = (OBJ, LIB, COMMON)
    try:
        OBJ = os.path.join(HERE, "build", name)
        os.makedirs(os.path.join(HERE, "variants"), exist_ok=True)
        LIB = os.path.join(HERE, "variants", "lib%s.so" % name)
        COMMON = COMMON + ["-D" + d for d in defines]
        return build(force=False)
    finally:
        OBJ, LIB, COMMON = saved


def build(force=False, verbose=False):
    os.makedirs(OBJ, exist_ok=True)
    hdrs = _headers()
    todo = [s for s in SOURCES
            if force or _newer(os.path.join(OBJ, s.replace(".cu", ".o")), [os.path.join(CSRC, s)] + hdrs)]
    if todo:
        with cf.ThreadPoolExecutor(max_workers=min(8, len(todo))) as ex:
            for src, log in ex.map(lambda s: _compile(s, verbose), todo):
                if verbose:
                    print("==", src)
                    print(log)
    objs = [os.path.join(OBJ, s.replace(".cu", ".o")) for s in SOURCES]
    if todo or _newer(LIB, objs):
        cmd = [NVCC, *ARCH, "-shared", "-o", LIB, *objs, "-Xcompiler", "-fPIC", "-ldl"]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError("link failed:\n%s\n%s" % (r.stdout, r.stderr))
    return LIB


if __name__ == "__main__":
    if "--variant" in sys.argv:
        i = sys.argv.index("--variant")
        print(build_variant(sys.argv[i + 1], sys.argv[i + 2:]))
    else:
        print(build(force="--force" in sys.argv, verbose="--verbose" in sys.argv))
