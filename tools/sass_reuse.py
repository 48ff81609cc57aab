"""sass_reuse.py -- how many DFMA of a kernel read three fresh 64-bit vector-register sources.
Measured on B200 (profiles/r2_microbench_dfma_operands.txt): a DFMA whose three sources are fresh vector registers
issues every 3 clk per scheduler; one with a source held in the operand reuse cache, in a uniform register or in the
constant bank every 2-2.2 clk.  Model of the reuse cache: one entry per source slot, filled by an operand that carries
.reuse, kept until another register is read in that slot, the register is overwritten, or control flow leaves the
block.  usage: tools/sass_fn.sh <obj> <kernel substring> | python tools/sass_reuse.py"""
import re
import sys

lines = [l for l in sys.stdin if re.search(r"/\*[0-9a-f]{4,5}\*/", l)]
cache, hist, ur = {}, {}, 0
for l in lines:
    m = re.search(r"\*/\s+(@!?U?P\d+\s+)?([A-Z0-9_.a-z]+)\s+(.*?);", l)
    if not m:
        cache = {}
        continue
    op, args = m.group(2), [a.strip() for a in m.group(3).split(",")]
    if op.startswith(("BRA", "BAR", "BSYNC", "CALL", "RET", "EXIT", "BRX")):
        cache = {}
    fresh, new = 0, dict(cache)
    for slot, a in enumerate(args[1:]):
        reg = re.sub(r"^[-|~!]*", "", a)
        flagged = reg.endswith(".reuse")
        reg = reg.replace(".reuse", "").rstrip("|")
        if re.fullmatch(r"R\d+", reg):
            if cache.get(slot) != reg:
                fresh += 1
            if flagged:
                new[slot] = reg
            else:
                new.pop(slot, None)
    md = re.fullmatch(r"R(\d+)", args[0] if args else "")
    if md:                                   # a (64-bit) write invalidates entries of the written pair
        d = int(md.group(1))
        for s_, r_ in list(new.items()):
            if int(r_[1:]) in (d, d + 1) or int(r_[1:]) + 1 == d:
                new.pop(s_)
    cache = new
    if op.startswith("DFMA"):
        hist[fresh] = hist.get(fresh, 0) + 1
        ur += bool(re.search(r"\bUR\d+|c\[", m.group(3)))
tot = sum(hist.values())
print("DFMA by fresh vector-register sources:", dict(sorted(hist.items())), "total", tot, "| with a uniform-register or constant source:", ur)
if tot:
    est = sum((3.0 if k >= 3 else 2.17) * v for k, v in hist.items()) / tot
    print("estimated issue cost %.2f clk per DFMA (all <= 2 fresh: 2.17)" % est)
