import csv, re, sys, os
sass_csv, dis_txt, kname, fname, line = sys.argv[1:6]; line = int(line)
addr2line = {}; insec = False; cur = None
for ln in open(dis_txt, errors="replace"):
    if ln.startswith("\t.section\t.text."):
        insec = kname in ln; cur = None; continue
    if not insec: continue
    m = re.match(r'\s*//## File "([^"]+)", line (\d+)(.*)', ln)
    if m: cur = (os.path.basename(m.group(1)), int(m.group(2))); continue
    m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(\S.*)", ln)
    if m and cur is not None: addr2line[int(m.group(1), 16)] = cur
rows = list(csv.reader(open(sass_csv)))
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address"); hdr = rows[hi]
ia, ie, it, isrc = hdr.index("Address"), hdr.index("Instructions Executed"), hdr.index("Thread Instructions Executed"), hdr.index("Source")
base = None; n = 0
for r in rows[hi + 1:]:
    try: a = int(r[ia], 16) if not r[ia].isdigit() else int(r[ia])
    except Exception: continue
    if base is None: base = a
    if addr2line.get(a - base) == (fname, line):
        n += 1
        if n <= int(sys.argv[6]) if len(sys.argv) > 6 else 60: print(hex(a - base), r[ie], r[it], r[isrc][:90])
print("count", n)
