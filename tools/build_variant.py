"""Experiment builds of libmibmi.so: tools/build_variant.py NAME [-DFOO=1 ...] -> tools/variants/NAME/libmibmi.so
(select at run time with MIBMI_LIB_PATH).  Only mibmi_pipe.cu and mibmi_tc.cu are recompiled; the other objects come from the
regular in-tree build."""
import os, subprocess, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mibminet_b200 import _build
name, flags = sys.argv[1], sys.argv[2:]
_build.build_library()
out = os.path.join(_build.ROOT, "tools", "variants", name)
os.makedirs(out, exist_ok=True)
nvcc = _build._nvcc()
VAR = ["mibmi_pipe.cu", "mibmi_tc.cu", "mibmi_tcr.cu", "mibmi_eegnet.cu"]
objs = [os.path.join(_build.LIB_DIR, s.replace(".cu", ".o")) for s in _build.SOURCES if s not in VAR]
procs = []
for src in VAR:
    o = os.path.join(out, src.replace(".cu", ".o"))
    procs.append(subprocess.Popen([nvcc, *_build.NVCC_FLAGS, *flags, "-c", os.path.join(_build.CSRC, src), "-o", o]))
    objs.append(o)
for pr in procs:
    assert pr.wait() == 0
subprocess.run([nvcc, "-shared", "-o", os.path.join(out, "libmibmi.so"), *objs, "-gencode", "arch=compute_100a,code=sm_100a",
                "-Xcompiler", "-fPIC"], check=True)
print(os.path.join(out, "libmibmi.so"))
