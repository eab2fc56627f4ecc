"""Build the host-fiber debug build of the kernels (tests only; see emu_runtime.h)."""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
CSRC = os.path.join(ROOT, "msgat_b200", "csrc")
OUT = os.path.join(HERE, "libmsgat_emu.so")


def sources():
    return sorted(os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith(".cu") and not f.startswith("tc_"))


def build(force=False):
    srcs = sources()
    deps = srcs + [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".h", ".cuh"))]
    deps += [os.path.join(HERE, "emu_runtime.h"), os.path.join(ROOT, "include", "msgat_sm100.h")]
    if not force and os.path.exists(OUT) and all(os.path.getmtime(OUT) >= os.path.getmtime(d) for d in deps):
        return OUT
    objs = []
    procs = []
    for s in srcs:
        o = os.path.join(HERE, os.path.basename(s)[:-3] + ".emu.o")
        objs.append(o)
        cmd = ["g++", "-x", "c++", "-std=c++17", "-O1", "-g", "-fPIC", "-DMSGAT_EMU", "-I", HERE, "-I", CSRC,
               "-Wno-unused-function", "-c", s, "-o", o]
        procs.append((subprocess.Popen(cmd), cmd))
    for pr, cmd in procs:
        if pr.wait() != 0:
            raise RuntimeError("emu build failed: " + " ".join(cmd))
    subprocess.check_call(["g++", "-shared", "-o", OUT] + objs)
    return OUT


if __name__ == "__main__":
    print(build(force="-f" in sys.argv))
