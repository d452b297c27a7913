"""Registers / stack / local memory of every kernel in the built device library, grouped by kernel family
(``cuobjdump --dump-resource-usage``; runs without a GPU).

    python scripts/kernel_resources.py                 # table on stdout
    python scripts/kernel_resources.py --out profiles/r2_kernel_resources.txt

A kernel that keeps arrays on the stack (STACK > a few dozen bytes of spill) or uses local memory pays for it in
every load / store of its inner loop: round 1's generic two-shot all-reduce and pull reduce-scatter did (312 / 608 B),
which is what the group-size-specialised kernels of round 2 removed. ``tests/test_kernel_resources.py`` pins the
kernels of record."""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SO = os.path.join(ROOT, "collective_communication_mpi_b200", "lib", "libccmpi_b200.so")


def demangle(names):
    out = subprocess.run(["c++filt"], input="\n".join(names), capture_output=True, text=True).stdout.splitlines()
    return dict(zip(names, out))


def collect(so=SO):
    txt = subprocess.run(["cuobjdump", "--dump-resource-usage", so], capture_output=True, text=True, check=True).stdout
    rows, cur = [], None
    for line in txt.splitlines():
        m = re.match(r"\s*Function (\S+):", line)
        if m:
            cur = m.group(1)
            continue
        m = re.match(r"\s*REG:(\d+) STACK:(\d+) SHARED:(\d+) LOCAL:(\d+)", line)
        if m and cur:
            rows.append((cur, *map(int, m.groups())))
            cur = None
    names = demangle([r[0] for r in rows])
    return [dict(mangled=r[0], name=names.get(r[0], r[0]), reg=r[1], stack=r[2], shared=r[3], local=r[4]) for r in rows]


def family(name: str) -> str:
    """Kernel name with its policy / variant, without the remaining template arguments."""
    name = name.replace("(anonymous namespace)::", "")
    m = re.match(r"(?:void )?([\w:]+)<(.*)>\(", name)
    if not m:
        return re.sub(r"\(.*", "", name.replace("void ", ""))
    base, args = m.groups()
    pol = re.search(r"(\w+Policy)", args)
    return f"{base}<{pol.group(1)}>" if pol else base


def main():
    rows = collect()
    fams = collections.defaultdict(list)
    for r in rows:
        fams[family(r["name"])].append(r)
    lines = [f"# {len(rows)} kernels in {os.path.relpath(SO, ROOT)} (cuobjdump --dump-resource-usage), by family",
             f"{'family':58s} {'kernels':>7s} {'regs min-max':>13s} {'stack max':>9s} {'local max':>9s}"]
    for f in sorted(fams):
        v = fams[f]
        lines.append(f"{f:58s} {len(v):7d} {min(x['reg'] for x in v):5d}-{max(x['reg'] for x in v):<5d}   "
                     f"{max(x['stack'] for x in v):9d} {max(x['local'] for x in v):9d}")
    text = "\n".join(lines) + "\n"
    if "--out" in sys.argv:
        open(sys.argv[sys.argv.index("--out") + 1], "w").write(text)
    sys.stdout.write(text)


if __name__ == "__main__":
    main()
