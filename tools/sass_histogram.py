"""Per-kernel histogram of the Blackwell-specific SASS mnemonics in libb200svs.so (cuobjdump -sass):
UTCHMMA (tcgen05.mma), UTCBAR (tcgen05.commit), LDTM / STTM (tcgen05.ld / st), UTMALDG / UTMASTG (TMA load / store), UTCATOMSWS
(TMEM alloc), SYNCS (mbarrier), UBLKCP etc.   python tools/sass_histogram.py > profiles/sass_mnemonics_r2.txt"""
import collections, os, re, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = os.path.join(ROOT, "music_generation_b200", "lib", "libb200svs.so")
sass = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
pat = re.compile(r"^\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)")
want = ("UTCHMMA", "UTCBAR", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UTMAPF", "UTCATOMSWS", "SYNCS", "UTMACMDFLUSH", "ELECT", "F2FP", "MUFU.EX2", "FADD2", "FMUL2", "HMMA", "ACQBULK", "UCGABAR")
per = collections.OrderedDict()
cur = None
for ln in sass.splitlines():
    m = re.match(r"\s*Function : (\S+)", ln)
    if m:
        cur = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
        cur = re.sub(r"\(.*", "", cur)
        per[cur] = collections.Counter()
        continue
    m = pat.match(ln)
    if m and cur:
        op = m.group(1)
        per[cur]["_total"] += 1
        for w in want:
            if op.startswith(w):
                key = w if not (w == "UTCHMMA" and ".2CTA" in op) else "UTCHMMA.2CTA"
                if w in ("UTCBAR",) and ".2CTA" in op:
                    key = "UTCBAR.2CTA"
                per[cur][key] += 1
arch = subprocess.run(["cuobjdump", "-lelf", lib], capture_output=True, text=True).stdout.strip().splitlines()
print("# SASS mnemonic histogram of music_generation_b200/lib/libb200svs.so (tools/sass_histogram.py)")
print("# ELF images:", ", ".join(a.split()[-1] for a in arch))
for fn, c in per.items():
    items = " ".join(f"{k}={v}" for k, v in sorted(c.items()) if k != "_total")
    print(f"{fn}: instructions={c['_total']} {items}")
tot = collections.Counter()
for c in per.values():
    tot.update(c)
print("TOTAL:", " ".join(f"{k}={v}" for k, v in sorted(tot.items())))
