"""Top instructions by stall samples of an `ncu -i rep --page source --csv` dump.  usage: python tools/ncu_source_top.py source.csv"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
hdr = rows[1]; data = rows[2:]
ix = {h:i for i,h in enumerate(hdr)}
def f(r,k):
    try: return float(r[ix[k]].replace(',',''))
    except: return 0.0
tot = sum(f(r,'# Samples') for r in data)
totI = sum(f(r,'Instructions Executed') for r in data)
totT = sum(f(r,'Thread Instructions Executed') for r in data)
print('instrs', len(data), 'samples', tot, 'warp-inst', totI, 'lanes', totT/totI)
# cumulative listing with region markers
top = sorted(range(len(data)), key=lambda i: -f(data[i],'# Samples'))[:45]
for i in sorted(top):
    r = data[i]
    print('%4d %-60s smp %5.1f%% exec %4.1f%% thr %4.1f  longsb %5d wait %4d math %4d br %4d notsel %4d short %4d lg %4d' % (i, r[ix['Source']].strip()[:60], 100*f(r,'# Samples')/tot, 100*f(r,'Instructions Executed')/totI, f(r,'Avg. Threads Executed'), f(r,'stall_long_sb'), f(r,'stall_wait'), f(r,'stall_math'), f(r,'stall_branch_resolving'), f(r,'stall_not_selected'), f(r,'stall_short_sb'), f(r,'stall_lg')))
