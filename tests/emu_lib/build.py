"""TEST INFRASTRUCTURE ONLY.  Builds tests/_build/libpdlp_b200_emu.so: every source of or-tools_b200/csrc compiled
with g++ as host C++ (-DPDLP_EMULATE) against tests/emu_lib/include and tests/kernel_emu/cuda_emu.h, kernel launches
rewritten by rewrite_launches.py.  See tests/emu_lib/include/cuda_runtime.h for what this is and is not."""
import concurrent.futures
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
CSRC = os.path.join(ROOT, "or-tools_b200", "csrc")
OUT = os.path.join(ROOT, "tests", "_build", "emu_lib")
LIB = os.path.join(ROOT, "tests", "_build", "libpdlp_b200_emu.so")
FAKE_NCCL = os.path.join(ROOT, "tests", "_build", "libfake_nccl.so")

sys.path.insert(0, HERE)
import rewrite_launches  # noqa: E402


def sources():
    names = sorted(os.listdir(CSRC))
    return [n for n in names if n.endswith((".cu", ".cc"))], [n for n in names if n.endswith((".cuh", ".h"))]


def stale(lib=LIB):
    if not os.path.exists(lib):
        return True
    t = os.path.getmtime(lib)
    deps = [os.path.join(CSRC, n) for n in os.listdir(CSRC)] + [os.path.join(HERE, "include", "cuda_runtime.h"), os.path.join(HERE, "rewrite_launches.py"),
                                                               os.path.join(ROOT, "tests", "kernel_emu", "cuda_emu.h"), os.path.abspath(__file__)]
    for base in (os.path.join(ROOT, "include"), os.path.join(ROOT, "include", "pdlp_b200")):
        deps += [os.path.join(base, n) for n in os.listdir(base) if n.endswith(".h")]
    return any(os.path.getmtime(d) > t for d in deps)


def build_fake_nccl():
    """NCCL for ranks that are host threads of one process (fake_nccl.cc); named through PDLP_B200_NCCL_LIBRARY."""
    src = os.path.join(HERE, "fake_nccl.cc")
    if not os.path.exists(FAKE_NCCL) or os.path.getmtime(FAKE_NCCL) < os.path.getmtime(src):
        os.makedirs(os.path.dirname(FAKE_NCCL), exist_ok=True)
        subprocess.run(["g++", "-std=c++20", "-O1", "-g", "-fPIC", "-shared", "-pthread", src, "-o", FAKE_NCCL], check=True)
    return FAKE_NCCL


def build(verbose=False, asan=False, tsan=False):
    """tsan=True: under ThreadSanitizer with one host thread per CUDA thread (libpdlp_b200_emu_tsan.so), the stand-in for
    compute-sanitizer's racecheck.  asan=True: the same library under AddressSanitizer (tests/_build/libpdlp_b200_emu_asan.so), the stand-in for
    compute-sanitizer's memcheck: "device" memory is heap memory, so a kernel or a copy that leaves its buffer is reported."""
    lib = LIB[:-3] + "_asan.so" if asan else (LIB[:-3] + "_tsan.so" if tsan else LIB)
    if not stale(lib):
        return lib
    os.makedirs(OUT, exist_ok=True)
    units, headers = sources()
    for n in units + headers:
        text, _ = rewrite_launches.rewrite(open(os.path.join(CSRC, n)).read())
        dst = os.path.join(OUT, n if not n.endswith(".cu") else n[:-3] + ".cu.cc")
        if not os.path.exists(dst) or open(dst).read() != text:
            open(dst, "w").write(text)
    flags = ["-std=c++20", "-O1", "-g", "-fPIC", "-pthread", "-ffp-contract=off", "-DPDLP_EMULATE", "-DPDLP_EMU_FIBERS", "-w",
             "-include", os.path.join(HERE, "include", "cuda_runtime.h"),
             "-I" + os.path.join(HERE, "include"), "-I" + os.path.join(ROOT, "tests", "kernel_emu"), "-I" + OUT,
             "-I" + os.path.join(ROOT, "include"), "-I" + os.path.join(ROOT, "include", "pdlp_b200")]
    if asan:
        flags += ["-fsanitize=address", "-fno-omit-frame-pointer", "-DPDLP_EMU_UCONTEXT"]
    if tsan:  # the racecheck stand-in: one host thread per CUDA thread (no fibers), so ThreadSanitizer sees every access pair
        flags = [f for f in flags if f != "-DPDLP_EMU_FIBERS"] + ["-fsanitize=thread"]
    suffix = ".asan.o" if asan else (".tsan.o" if tsan else ".o")

    def compile_one(n):
        src = os.path.join(OUT, n if not n.endswith(".cu") else n[:-3] + ".cu.cc")
        obj = src + suffix
        run = subprocess.run(["g++"] + flags + ["-c", src, "-o", obj], capture_output=True, text=True)
        return n, obj, run

    objs = []
    with concurrent.futures.ThreadPoolExecutor(max_workers=os.cpu_count() or 4) as pool:
        for n, obj, run in pool.map(compile_one, units):
            if run.returncode != 0:
                raise RuntimeError(f"emulated build of {n} failed:\n{run.stderr[:6000]}")
            if verbose and run.stderr:
                print(run.stderr[:2000])
            objs.append(obj)
    run = subprocess.run(["g++", "-shared", "-pthread"] + (["-fsanitize=address"] if asan else []) + (["-fsanitize=thread"] if tsan else []) + ["-o", lib] + objs + ["-ldl"], capture_output=True, text=True)
    if run.returncode != 0:
        raise RuntimeError("link of the emulated library failed:\n" + run.stderr[:4000])
    return lib


if __name__ == "__main__":
    print(build(verbose=True, asan="--asan" in sys.argv, tsan="--tsan" in sys.argv))
    print(build_fake_nccl())
