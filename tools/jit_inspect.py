#!/usr/bin/env python
"""Development tool: fetches the generated kernel of a GERG model from libteqpb200 (no GPU needed), writes source / constants / cubin,
disassembles the cubin and prints the SASS opcode mix -- the per-state-point instruction budget of the per-model kernel.

    python tools/jit_inspect.py C3 0 4 [outdir]
"""
import collections
import ctypes as C
import json
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402

import teqp_b200  # noqa: E402
from teqp_b200.workloads import WORKLOADS  # noqa: E402


def generated(model, NT, ND, which):
    lib = teqp_b200.lib()
    n = C.c_ulonglong(0)
    err = C.create_string_buffer(4000)
    rc = lib.teqpb200_generated_kernel(model._uuid, NT, ND, which, None, C.byref(n), err, 4000)
    if rc:
        raise RuntimeError(err.value.decode())
    buf = C.create_string_buffer(n.value)
    rc = lib.teqpb200_generated_kernel(model._uuid, NT, ND, which, buf, C.byref(n), err, 4000)
    if rc:
        raise RuntimeError(err.value.decode())
    return buf.raw[:n.value]


def main():
    wl, NT, ND = sys.argv[1], int(sys.argv[2]), int(sys.argv[3])
    out = sys.argv[4] if len(sys.argv) > 4 else "/tmp/jit"
    os.makedirs(out, exist_ok=True)
    spec = WORKLOADS[wl]["spec"] if wl in WORKLOADS else json.loads(wl)
    m = teqp_b200.make_model(spec)
    src = generated(m, NT, ND, 0).rstrip(b"\0")
    open(os.path.join(out, "k.cu"), "wb").write(src)
    consts = np.frombuffer(generated(m, NT, ND, 1), dtype=np.float64)
    consts.tofile(os.path.join(out, "k.consts"))
    cubin = generated(m, NT, ND, 2)
    open(os.path.join(out, "k.cubin"), "wb").write(cubin)
    sass = subprocess.run(["cuobjdump", "-sass", os.path.join(out, "k.cubin")], stdout=subprocess.PIPE, text=True).stdout
    open(os.path.join(out, "k.sass"), "w").write(sass)
    res = subprocess.run(["cuobjdump", "-res-usage", os.path.join(out, "k.cubin")], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True).stdout
    ops = collections.Counter()
    for line in sass.splitlines():
        mm = re.match(r"\s+/\*[0-9a-f]+\*/\s+(@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
        if mm:
            ops[mm.group(2).split(".")[0]] += 1
    tot = sum(ops.values())
    fp64 = sum(v for k, v in ops.items() if k in ("DFMA", "DMUL", "DADD", "DSETP", "DMNMX"))
    print("source bytes", len(src), "consts", consts.size, "cubin bytes", len(cubin))
    print(res.strip().splitlines()[-1] if res.strip() else "")
    print("static SASS instructions: %d, FP64 %d (%.1f %%)" % (tot, fp64, 100.0 * fp64 / tot))
    print(", ".join("%s %d" % kv for kv in ops.most_common(24)))


if __name__ == "__main__":
    main()
