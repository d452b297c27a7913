"""stdin: `cuobjdump -sass -fun ...`; stdout: the function's instructions as `address  instruction`, without the
encoding words and the per-object headers (keeps the listings a tenth of the size)."""
import re
import sys

on = False
for l in sys.stdin:
    if "Function :" in l:
        on = True
    if not on or not l.strip() or re.match(r"^\s+/\* 0x[0-9a-f]+ \*/\s*$", l):
        continue
    l = re.sub(r"\s*/\* 0x[0-9a-f]+ \*/\s*$", "", l.rstrip("\n"))
    print(re.sub(r"^\s+/\*([0-9a-f]+)\*/\s+", r"\1  ", l))
