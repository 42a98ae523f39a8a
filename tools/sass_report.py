"""Static evidence for profiles/: per-kernel register / spill figures (nvcc -Xptxas -v) and SASS opcode
histograms (cuobjdump -sass) of csrc/libgradavg.so.  Runs on the CPU box (no GPU needed).
    python tools/sass_report.py > profiles/r02_sass_ptxas.md"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CSRC = os.path.join(ROOT, "pytorch-distributed-example_b200", "csrc")
WATCH = ["LDG", "STG", "LD", "ST", "LDL", "STL", "LDS", "STS", "LDGMC", "MEMBAR", "ATOM", "ATOMS", "RED", "BAR", "FADD", "FCHK", "CALL",
         "UTMALDG", "UTMASTG", "UBLKCP"]


def demangle(names):
    out = subprocess.run(["c++filt"], input="\n".join(names), capture_output=True, text=True).stdout.splitlines()
    return [re.sub(r"\(anonymous namespace\)::", "", o).split("(")[0] for o in out]


def ptxas():
    rows = []
    for unit in ("gradavg.cu", "gradavg_view.cu"):
        err = subprocess.run(["nvcc", "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17", "-Xcompiler", "-fPIC",
                              "-Xptxas", "-v", "-c", "-o", "/dev/null", unit], cwd=CSRC, capture_output=True, text=True).stderr
        cur = None
        for line in err.splitlines():
            m = re.search(r"Compiling entry function '(\S+)'", line)
            if m:
                cur = {"name": m.group(1)}
                rows.append(cur)
            m = re.search(r"(\d+) bytes stack frame, (\d+) bytes spill stores, (\d+) bytes spill loads", line)
            if m and cur is not None:
                cur["stack"], cur["spill_st"], cur["spill_ld"] = map(int, m.groups())
            m = re.search(r"Used (\d+) registers", line)
            if m and cur is not None:
                cur["regs"] = int(m.group(1))
    names = demangle([r["name"] for r in rows])
    for r, n in zip(rows, names):
        r["pretty"] = n
    return rows


def sass():
    txt = subprocess.run(["cuobjdump", "-sass", os.path.join(CSRC, "libgradavg.so")], capture_output=True, text=True).stdout
    hist, cur = collections.OrderedDict(), None
    for line in txt.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            cur = m.group(1)
            hist[cur] = collections.Counter()
            continue
        m = re.match(r"\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)", line)
        if m and cur:
            hist[cur][m.group(1)] += 1
            hist[cur]["__all__"] += 1
        if cur and "128" in line and re.search(r"\b(LDG|STG)\.E(\.\w+)*\.128", line):
            hist[cur]["__128bit__"] += 1
    return hist


def family(pretty):
    return re.sub(r"<.*", "", pretty)


def main():
    rows = ptxas()
    print("# Round 2: static evidence for `csrc/libgradavg.so` (`tools/sass_report.py`, nvcc 12.9, sm_100a)\n")
    print("## Registers and spills (`nvcc -Xptxas -v`), %d kernel instantiations\n" % len(rows))
    fams = collections.OrderedDict()
    for r in rows:
        fams.setdefault(family(r["pretty"]), []).append(r)
    print("| kernel family | instantiations | registers (min-max) | stack frame bytes (max) | spill stores / loads (max) |")
    print("|---|---|---|---|---|")
    for f, rs in fams.items():
        print("| `%s` | %d | %d-%d | %d | %d / %d |" % (f, len(rs), min(r["regs"] for r in rs), max(r["regs"] for r in rs),
                                                   max(r.get("stack", 0) for r in rs), max(r.get("spill_st", 0) for r in rs),
                                                   max(r.get("spill_ld", 0) for r in rs)))
    spilling = [r for r in rows if r.get("spill_st", 0) or r.get("spill_ld", 0)]
    print("\nInstantiations with spills: **%d**%s\n" % (len(spilling), "".join("\n* `%s`: %d / %d bytes" % (r["pretty"], r["spill_st"], r["spill_ld"]) for r in spilling)))
    hist = sass()
    pretty = dict(zip(hist.keys(), demangle(list(hist.keys()))))
    agg = collections.OrderedDict()
    for k, c in hist.items():
        a = agg.setdefault(family(pretty[k]), collections.Counter())
        a.update(c)
        a["__n__"] += 1
    print("## SASS opcode counts per kernel family (`cuobjdump -sass`, summed over the family's instantiations)\n")
    print("| family | n | instructions | " + " | ".join(WATCH) + " | 128-bit LDG/STG |")
    print("|---|---|---|" + "---|" * (len(WATCH) + 1))
    tot = collections.Counter()
    for f, c in agg.items():
        tot.update(c)
        print("| `%s` | %d | %d | " % (f, c["__n__"], c["__all__"]) + " | ".join(str(c.get(w, 0)) for w in WATCH) + " | %d |" % c["__128bit__"])
    print("| **all** | %d | %d | " % (tot["__n__"], tot["__all__"]) + " | ".join(str(tot.get(w, 0)) for w in WATCH) + " | %d |" % tot["__128bit__"])
    print("""
Reading: `LDL`/`STL` = local memory (spills, dynamically indexed arrays), `LD`/`ST` = generic-space accesses whose
state space the compiler could not prove, `LDGMC` = `multimem.ld_reduce` (NVSwitch in-switch reduction; `multimem.st` assembles to a
plain `STG...SYS` on a multicast address), `MEMBAR` = the system-scope fences of the flag protocol, `FCHK`/`CALL` = the slow path of the
IEEE division `__fdiv_rn` (taken for denormal / huge operands only).  No `UTMALDG`/`UTMASTG`/`UBLKCP`: TMA bulk copies were measured
against 16-byte SIMT stores for this traffic (profiles/r02_nvlink_probe.md: 649 vs 630 GB/s all-to-all) and not adopted; north_star
specifies 16-byte SIMT vectors.""")


if __name__ == "__main__":
    sys.exit(main())
