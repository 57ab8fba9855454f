"""Build libede_sm100.so in-tree with nvcc (sm_100a only; cross-compiles without a GPU)."""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB_DIR = os.path.join(HERE, "lib")
LIB = os.path.join(LIB_DIR, "libede_sm100.so")
SOURCES = ["api.cu", "conv_direct.cu", "conv_umma.cu", "bandwidth.cu", "dwconv.cu", "dwconv_tile.cu", "upconv.cu", "stem.cu", "sobel.cu", "preprocess.cu", "prepack.cu", "render.cu"]


def _stale() -> bool:
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    deps = [os.path.join(CSRC, f) for f in os.listdir(CSRC)] + [os.path.join(HERE, "..", "include", "ede.h")]
    return any(os.path.getmtime(d) > t for d in deps)


def _compile(args):
    nvcc, src, obj, verbose = args
    cmd = [nvcc, "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
           "-Xcompiler", "-fPIC", "-c", "-o", obj, src]
    if verbose:
        cmd.insert(1, "-Xptxas=-v")
    r = subprocess.run(cmd, capture_output=True, text=True)
    return r.returncode, r.stdout + r.stderr


def build(force: bool = False, verbose: bool = False) -> str:
    """One object per source (compiled in parallel, re-used while the source and the headers are older), then a
    shared-library link.  `force` recompiles everything."""
    if not force and not _stale():
        return LIB
    from concurrent.futures import ThreadPoolExecutor
    obj_dir = os.path.join(LIB_DIR, "obj")
    os.makedirs(obj_dir, exist_ok=True)
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    headers = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cuh", ".h"))]
    headers.append(os.path.join(HERE, "..", "include", "ede.h"))
    t_hdr = max(os.path.getmtime(h) for h in headers)
    jobs, objs = [], []
    for s in SOURCES:
        src, obj = os.path.join(CSRC, s), os.path.join(obj_dir, s + ".o")
        objs.append(obj)
        if force or not os.path.exists(obj) or os.path.getmtime(obj) < max(os.path.getmtime(src), t_hdr):
            jobs.append((nvcc, src, obj, verbose))
    with ThreadPoolExecutor(max_workers=max(1, min(len(jobs), os.cpu_count() or 4))) as ex:
        for rc, out in ex.map(_compile, jobs):
            if rc != 0:
                raise RuntimeError("nvcc failed:\n" + out)
            if verbose:
                print(out)
    r = subprocess.run([nvcc, "-shared", "-o", LIB] + objs, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("nvcc link failed:\n" + r.stdout + r.stderr)
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
