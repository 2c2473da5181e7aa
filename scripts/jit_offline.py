"""Compiles the topology-specialised kernel OFFLINE (NVRTC via cuda-python, no GPU needed) for a saved prelude, so that
register counts / SASS can be studied before spending GPU time.
usage: jit_offline.py prelude.txt out.cubin [-DNAME=VALUE ...]     (prelude: the TJ_* lines of an OBGPU_TINY_JIT_DUMP file)"""
import os
import re
import subprocess
import sys
import tempfile

from cuda.bindings import nvrtc

root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
csrc = os.path.join(root, "openbeam_b200", "csrc")
with tempfile.TemporaryDirectory() as td:
    subprocess.check_call([sys.executable, os.path.join(csrc, "gen_jit_src.py"), csrc, os.path.join(td, "x.cpp"), os.path.join(td, "raw.cu")])
    raw = open(os.path.join(td, "raw.cu")).read()
prelude = open(sys.argv[1]).read()
for d in sys.argv[3:]:
    m = re.match(r"-D(\w+)=(.*)", d)
    prelude = re.sub(r"#define %s .*" % m.group(1), "", prelude) + "#define %s %s\n" % (m.group(1), m.group(2))
src = (prelude + raw).encode()
err, prog = nvrtc.nvrtcCreateProgram(src, b"k_tiny_jit.cu", 0, [], [])
opts = [b"--gpu-architecture=sm_100a", b"--std=c++17", b"-lineinfo", b"-default-device"]
(err,) = nvrtc.nvrtcCompileProgram(prog, len(opts), opts)
_, n = nvrtc.nvrtcGetProgramLogSize(prog)
log = b" " * n
nvrtc.nvrtcGetProgramLog(prog, log)
txt = log.decode(errors="replace")
print("\n".join(l for l in txt.split("\n") if "error" in l or "#177" not in l)[:4000] if int(err) else "compiled")
if int(err):
    sys.exit(1)
_, sz = nvrtc.nvrtcGetCUBINSize(prog)
cubin = b" " * sz
nvrtc.nvrtcGetCUBIN(prog, cubin)
open(sys.argv[2], "wb").write(cubin)
print(subprocess.run(["cuobjdump", "-res-usage", sys.argv[2]], capture_output=True, text=True).stdout.strip().split("\n")[-1])
