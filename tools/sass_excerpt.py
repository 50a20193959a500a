"""SASS evidence per kernel, checkable without the .so:  python tools/sass_excerpt.py > profiles/r02_sass_excerpt.txt

For every kernel of libthormang_b200.so (cuobjdump -sass, sm_100a cubin): instruction count, registers (from
cuobjdump -res-usage), and the counts + first occurrences of the instructions that prove the design claims of DESIGN.md:
  UBLKCP   cp.async.bulk (1-D bulk copy through the TMA engine, global<->shared)
  UBLKPF   cp.async.bulk.prefetch.L2
  SYNCS    mbarrier arrive / expect_tx / try_wait
  ACQBULK / FENCE.VIEW.ASYNC  async-proxy fences around the bulk store
  LDG/STG.E.128, .ENL2.256     128/256-bit global accesses
  LDS/STS.64/.128             vector shared-memory accesses
  RED / REDG                  statistics partials
  SHFL                        warp shuffles
"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "thormang_isaacgym_b200", "libthormang_b200.so")
KEYS = ["UBLKCP", "UBLKPF", "SYNCS", "FENCE", "LDG", "STG", "LDS", "STS", "RED", "SHFL", "MUFU", "BAR", "ACQBULK", "ERRBAR"]
WIDE = re.compile(r"\.(64|128|256)\b")


def demangle(names):
    out = subprocess.run(["c++filt"], input="\n".join(names), capture_output=True, text=True).stdout.splitlines()
    return dict(zip(names, out))


def main():
    sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    res = subprocess.run(["cuobjdump", "-res-usage", LIB], capture_output=True, text=True).stdout
    regs = {}
    cur = None
    for line in res.splitlines():
        m = re.search(r"Function (\S+):", line)
        if m:
            cur = m.group(1)
        m = re.search(r"REG:(\d+).*?SHARED:(\d+)", line)
        if m and cur:
            regs[cur] = (int(m.group(1)), int(m.group(2)))
    kernels = collections.OrderedDict()
    cur = None
    for line in sass.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            cur = m.group(1)
            kernels[cur] = []
            continue
        m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(.*?);", line)
        if m and cur:
            kernels[cur].append((m.group(1), m.group(2).strip()))
    names = demangle(list(kernels))
    print(f"# cuobjdump -sass {os.path.relpath(LIB, ROOT)} (sm_100a); {len(kernels)} kernels")
    print("# per kernel: instruction count, registers, static shared bytes, opcode-family counts (wide = .64/.128/.256 variants),")
    print("# then the first two occurrences of the bulk-copy / mbarrier instructions")
    for mangled, ins in kernels.items():
        name = names.get(mangled, mangled)
        if len(name) > 140:
            name = name[:137] + "..."
        cnt = collections.Counter()
        wide = collections.Counter()
        firsts = collections.defaultdict(list)
        for addr, text in ins:
            op = text.split()[1] if text.startswith("@") and len(text.split()) > 1 else text.split()[0]
            fam = op.split(".")[0]
            for k in KEYS:
                if fam.startswith(k):
                    cnt[k] += 1
                    w = WIDE.search(op)
                    if w and k in ("LDG", "STG", "LDS", "STS"):
                        wide[f"{k}.{w.group(1)}"] += 1
                    if k in ("UBLKCP", "UBLKPF", "SYNCS") and len(firsts[k]) < 2:
                        firsts[k].append(f"    /*{addr}*/ {text}")
                    break
        r = regs.get(mangled)
        rtxt = f"regs={r[0]} smem_static={r[1]}" if r else ""
        print(f"\n{name}\n  instructions={len(ins)} {rtxt}")
        print("  " + " ".join(f"{k}={cnt[k]}" for k in KEYS if cnt[k]))
        if wide:
            print("  wide: " + " ".join(f"{k}={v}" for k, v in sorted(wide.items())))
        for k in ("UBLKCP", "UBLKPF", "SYNCS"):
            for t in firsts[k]:
                print(t)


if __name__ == "__main__":
    sys.exit(main())
