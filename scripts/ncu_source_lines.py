"""Per-source-line instruction counts and stall samples from `ncu --page source --csv --print-source cuda,sass`."""
import csv, sys, collections
rows = list(csv.reader(open(sys.argv[1])))
topn = int(sys.argv[2]) if len(sys.argv) > 2 else 40
cur = None
hdr = None
agg = {}
for r in rows:
    if not r:
        continue
    if r[0] == 'File Path':
        cur = r[1].split('/')[-1]; continue
    if r[0] == 'Function Name':
        continue
    if r[0] == 'Line No':
        hdr = r; continue
    if r[0] == '' or cur is None:
        continue
    try:
        line = int(r[0])
    except ValueError:
        continue
    # the source text may contain commas -> columns shift; take numeric fields from the right using header length
    ncol = len(hdr)
    extra = len(r) - ncol
    vals = r[2 + extra:]
    h = hdr[2:]
    d = dict(zip(h, vals))
    def num(k):
        try: return float(d.get(k, '0').replace(',', ''))
        except ValueError: return 0.0
    src = ','.join(r[1:2 + extra])
    agg[(cur, line)] = (num('Instructions Executed'), num('# Samples'), num('stall_long_sb'), num('stall_wait'), num('stall_barrier'), num('stall_short_sb'), num('stall_no_inst'), num('stall_branch_resolving'), src.strip()[:90])
tot_i = sum(v[0] for v in agg.values()); tot_s = sum(v[1] for v in agg.values())
print("total inst %.4g samples %.0f" % (tot_i, tot_s))
print("by instructions:")
for (f, l), v in sorted(agg.items(), key=lambda kv: -kv[1][0])[:topn]:
    print("%-16s %5d  inst %5.2f%%  samp %5.2f%%  lsb %4.0f wait %4.0f bar %4.0f ssb %4.0f | %s" % (f, l, 100*v[0]/tot_i, 100*v[1]/tot_s, v[2], v[3], v[4], v[5], v[8]))
print("by samples:")
for (f, l), v in sorted(agg.items(), key=lambda kv: -kv[1][1])[:topn]:
    print("%-16s %5d  inst %5.2f%%  samp %5.2f%%  lsb %4.0f wait %4.0f bar %4.0f ssb %4.0f noi %4.0f br %4.0f | %s" % (f, l, 100*v[0]/tot_i, 100*v[1]/tot_s, v[2], v[3], v[4], v[5], v[6], v[7], v[8]))
# regions of wfa_core.cuh
import bisect
regions = [(1,251,'hdr'),(252,322,'lcp/window'),(323,443,'misc/extend0'),(444,566,'publish/step_finish'),(567,744,'wf_step scalar'),(745,848,'helpers'),(849,960,'wf_setup_dir'),(961,1097,'multi: prologue+claim'),(1098,1136,'multi: row loads+shuffles'),(1137,1178,'multi: fringe masks'),(1179,1206,'multi: recurrence+bounds'),(1207,1243,'multi: gap stores+trims'),(1244,1288,'multi: extension'),(1289,1313,'multi: M store+endcell'),(1314,1378,'multi: publish'),(1379,1394,'red_reset'),(1395,1515,'overlap'),(1516,1626,'find_breakpoint'),(1627,1760,'backtrace'),(1761,1800,'wf_base'),(1801,2000,'driver')]
ra = collections.defaultdict(lambda: [0,0])
for (f, l), v in agg.items():
    if f != 'wfa_core.cuh':
        ra[f][0] += v[0]; ra[f][1] += v[1]; continue
    for a, b, n in regions:
        if a <= l <= b:
            ra[n][0] += v[0]; ra[n][1] += v[1]; break
print("regions:")
for n, v in sorted(ra.items(), key=lambda kv: -kv[1][1]):
    print("  %-28s inst %5.1f%%  samples %5.1f%%" % (n, 100*v[0]/tot_i, 100*v[1]/tot_s))
