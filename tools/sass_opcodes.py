#!/usr/bin/env python
"""Static SASS opcode table of the frame kernels of a built libsr_cuda.so (cuobjdump -sass): total instructions and the
ones that matter for the "FP64 on the hot path" question: D* arithmetic, 64-bit conversions, MUFU,
integer multiply-add, shared / global memory.   python tools/sass_opcodes.py [lib.so] > profiles/r02_sass_opcodes.md"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "software_renderer_b200", "libsr_cuda.so")
sass = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True, check=True).stdout
demangle = lambda n: subprocess.run(["c++filt", n], capture_output=True, text=True).stdout.strip().split("(")[0].replace("void ", "").replace("sr::", "")
kern, ops = None, collections.OrderedDict()
for line in sass.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        kern = demangle(m.group(1))
        ops[kern] = collections.Counter()
        continue
    m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
    if m and kern:
        ops[kern][m.group(1)] += 1
groups = [("FP64 arith (DMUL/DFMA/DADD/DSETP)", lambda o: re.match(r"D(MUL|FMA|ADD|SETP|MNMX)", o)),
          ("F2I/I2F/F2F with .F64 or .64", lambda o: re.match(r"(F2I|I2F|F2F|I2FP)", o) and ("F64" in o or ".64" in o)),
          ("F2I/I2F 32-bit", lambda o: re.match(r"(F2I|I2F|I2FP|F2IP)", o) and "F64" not in o and ".64" not in o),
          ("MUFU (rcp, ...)", lambda o: o.startswith("MUFU")), ("FMUL/FADD/FFMA", lambda o: re.match(r"F(MUL|ADD|FMA)", o)),
          ("IMAD/IADD3/LEA", lambda o: re.match(r"(IMAD|IADD3|LEA|IADD)", o)), ("LDS/STS", lambda o: re.match(r"(LDS|STS)", o)),
          ("LDG/STG/ATOMG/RED", lambda o: re.match(r"(LDG|STG|ATOMG|REDG|ATOM|RED)", o)), ("BAR/WARPSYNC/SHFL/VOTE", lambda o: re.match(r"(BAR|WARPSYNC|SHFL|VOTE|MATCH|REDUX)", o))]
print(f"# Static SASS opcode counts per kernel — `{os.path.relpath(lib, ROOT)}` (cuobjdump -sass; instructions in the binary, not executed)\n")
print("| kernel | total | " + " | ".join(g for g, _ in groups) + " |")
print("|---|---:|" + "---:|" * len(groups))
for k, c in ops.items():
    if not re.match(r"k_(tile|setup|transform|resolve|late_boxes|points|na_)", k):
        continue
    print(f"| `{k}` | {sum(c.values())} | " + " | ".join(str(sum(n for o, n in c.items() if f(o))) for _, f in groups) + " |")
