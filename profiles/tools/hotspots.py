"""Join an ncu SASS-page CSV (ncu -i X.ncu-rep --page source --csv --print-source sass) with nvdisasm
--print-line-info output of the same cubin, and aggregate executed warp instructions + stall samples per
source function / line.  Usage: hotspots.py sass.csv dis.txt kernel_substr [csrc_dir]"""
import csv, re, sys, os, collections

sass_csv, dis_txt, kname = sys.argv[1:4]
csrc = sys.argv[4] if len(sys.argv) > 4 else None

# ---- nvdisasm: address -> (file, line) for the kernel's text section
addr2line = {}
insec = False
cur = None
for ln in open(dis_txt, errors="replace"):
    if ln.startswith("\t.section\t.text."):
        insec = kname in ln
        cur = None
        continue
    if not insec:
        continue
    m = re.match(r'\s*//## File "([^"]+)", line (\d+)(.*)', ln)
    if m:
        if "inlined at" in m.group(3) and cur is not None and False:
            pass
        cur = (os.path.basename(m.group(1)), int(m.group(2)))
        continue
    m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(\S.*)", ln)
    if m and cur is not None:
        addr2line[int(m.group(1), 16)] = cur

# ---- function ranges from the sources
ranges = collections.defaultdict(list)
if csrc:
    for f in os.listdir(csrc):
        lines = open(os.path.join(csrc, f), errors="replace").read().split("\n")
        for i, l in enumerate(lines):
            m = re.match(r"^(?:template.*>\s*)?(?:G2_HDN?|static|__global__|inline)\b.*?\b([A-Za-z_0-9]+)\s*\(", l)
            if m and not l.strip().endswith(";"):
                ranges[f].append((i + 1, m.group(1)))

def func_of(f, line):
    best = "?"
    for (start, name) in ranges.get(f, []):
        if start <= line:
            best = name
        else:
            break
    return best

rows = list(csv.reader(open(sass_csv)))
hdr_i = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr = rows[hdr_i]
ia, ie, is_, it = hdr.index("Address"), hdr.index("Instructions Executed"), hdr.index("# Samples"), hdr.index("Thread Instructions Executed")
base = None
byfunc = collections.Counter(); sfunc = collections.Counter(); tfunc = collections.Counter()
byline = collections.Counter(); sline = collections.Counter()
tot = 0; stot = 0
for r in rows[hdr_i + 1:]:
    if len(r) <= max(ia, ie, is_):
        continue
    try:
        a = int(r[ia], 16) if not r[ia].isdigit() else int(r[ia])
    except ValueError:
        continue
    if base is None:
        base = a
    off = a - base
    ex = int(r[ie] or 0); sm = int(r[is_] or 0); th = int(r[it] or 0)
    f, l = addr2line.get(off, ("?", 0))
    fn = func_of(f, l)
    byfunc[fn] += ex; sfunc[fn] += sm; tfunc[fn] += th
    byline[(f, l)] += ex; sline[(f, l)] += sm
    tot += ex; stot += sm
print(f"total warp instr {tot}  samples {stot}")
print(f"{'function':32s} {'instr%':>7s} {'stall%':>7s} {'lanes':>6s}")
for fn, ex in sfunc.most_common(40):
    print(f"{fn:32s} {100*byfunc[fn]/max(tot,1):7.2f} {100*sfunc[fn]/max(stot,1):7.2f} {tfunc[fn]/max(byfunc[fn],1):6.1f}")
print("\ntop lines by stall samples")
for (f, l), s in sline.most_common(40):
    print(f"{f}:{l:5d} instr% {100*byline[(f,l)]/max(tot,1):6.2f} stall% {100*s/max(stot,1):6.2f}")
if os.environ.get("HOTSPOTS_BY_INSTR"):
    tl = collections.Counter()
    for r in rows[hdr_i + 1:]:
        pass
    print("\ntop lines by issued warp instructions")
    for (f, l), ex in byline.most_common(int(os.environ["HOTSPOTS_BY_INSTR"])):
        print(f"{f}:{l:5d} instr% {100*ex/max(tot,1):6.2f} stall% {100*sline[(f,l)]/max(stot,1):6.2f}")
if os.environ.get("HOTSPOTS_MEM"):
    isrc = hdr.index("Source")
    ld = collections.Counter(); st = collections.Counter(); gl = collections.Counter()
    for r in rows[hdr_i + 1:]:
        if len(r) <= max(ia, ie, isrc):
            continue
        try:
            a = int(r[ia], 16) if not r[ia].isdigit() else int(r[ia])
        except ValueError:
            continue
        f, l = addr2line.get(a - base, ("?", 0))
        fn = func_of(f, l)
        ex = int(r[ie] or 0)
        op = r[isrc]
        w = 4 if ".128" in op else 2 if ".64" in op else 1
        if "LDL" in op: ld[fn] += ex * w
        elif "STL" in op: st[fn] += ex * w
        elif "LDG" in op or "STG" in op or " LD." in op or " ST." in op: gl[fn] += ex * w
    tl, ts, tg = sum(ld.values()), sum(st.values()), sum(gl.values())
    print(f"\nlocal-memory warp-level word accesses: loads {tl} stores {ts}; global {tg}")
    for fn, n in (ld + st).most_common(int(os.environ["HOTSPOTS_MEM"])):
        print(f"{fn:32s} ld% {100*ld[fn]/max(tl,1):6.2f} st% {100*st[fn]/max(ts,1):6.2f} global% {100*gl[fn]/max(tg,1):6.2f}")
