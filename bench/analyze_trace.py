"""Summarises gpurun_out/trace.json.gz (bench/timeline.py): per-stream busy time, gaps on the main
chain, kernels by in-situ duration for the LAST captured step."""
import collections, gzip, json, re, sys
tr = json.load(gzip.open(sys.argv[1] if len(sys.argv) > 1 else 'gpurun_out/trace.json.gz'))
ev = [e for e in tr['traceEvents'] if e.get('cat') in ('kernel', 'gpu_memcpy', 'gpu_memset')]
ev.sort(key=lambda e: e['ts'])
# split into steps at the imagenet_prep kernel
starts = [i for i, e in enumerate(ev) if 'imagenet_prep' in e['name']]
lo = starts[-1]
step = ev[lo:]
t0 = step[0]['ts']; t1 = max(e['ts'] + e['dur'] for e in step)
print(f"last step: {len(step)} events, span {(t1 - t0) / 1000:.3f} ms")
by_stream = collections.defaultdict(list)
for e in step: by_stream[e['args']['stream']].append(e)
for s, es in by_stream.items():
    busy = sum(e['dur'] for e in es)
    print(f"stream {s}: {len(es)} events busy {busy / 1000:.3f} ms, first {es[0]['name'][:40]}")
main = max(by_stream.values(), key=len)
gaps = []
for a, b in zip(main, main[1:]):
    gaps.append((b['ts'] - (a['ts'] + a['dur']), a['name'][:38], b['name'][:38]))
tot_gap = sum(g for g, _, _ in gaps if g > 0)
print(f"main chain: busy {sum(e['dur'] for e in main) / 1000:.3f} ms, positive gaps {tot_gap / 1000:.3f} ms over {len(gaps)} boundaries, median {sorted(g for g,_,_ in gaps)[len(gaps)//2]:.2f} us")
for g in sorted(gaps, reverse=True)[:8]: print("   gap %.1f us  %s -> %s" % g)
def short(n):
    n = re.sub(r'\(.*', '', n); return re.sub(r'void |drn::|\(anonymous namespace\)::|at::native::', '', n)[:60]
agg = collections.defaultdict(lambda: [0, 0.0])
for e in step:
    a = agg[(e['args']['stream'], short(e['name']))]; a[0] += 1; a[1] += e['dur']
for (s, n), (c, d) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:16]:
    print(f"{d:9.1f} us n={c:3d} avg {d / c:7.1f} stream {s} {n}")
if len(sys.argv) > 2:
    for e in main:
        if sys.argv[2] in e['name']: print(f"{e['dur']:8.1f} {short(e['name'])}")
