"""Per-kernel table of the LAST training step in a chrome trace written by bench/timeline.py:
python bench/step_table.py gpurun_out/timeline_cifar10_16.json[.gz] [first-kernel-substring]"""
import collections, gzip, json, re, sys
path = sys.argv[1]
tr = json.load(gzip.open(path) if path.endswith(".gz") else open(path))
ev = sorted((e for e in tr["traceEvents"] if e.get("cat") in ("kernel", "gpu_memcpy", "gpu_memset")), key=lambda e: e["ts"])
first = sys.argv[2] if len(sys.argv) > 2 else "augment"
starts = [i for i, e in enumerate(ev) if first in e["name"] or "imagenet_prep" in e["name"]]
step = ev[starts[-1]:]
def short(n):
    m = re.search(r"\d+(conv_igemm_kernel|conv_wgrad_kernel|\w+?_kernel)", n)
    base = m.group(1) if m else n[:40]
    t = re.search(r"ILi(\d+)ELi(\d+)ELb(\d)ELb(\d)ELi(-?\d+)E", n)
    if t: base += f"<{t.group(1)},{t.group(2)},pre{t.group(3)},two{t.group(4)},epi{t.group(5)}>"
    t = re.search(r"wgrad_kernelILi(\d+)ELi(\d+)ELb(\d)", n)
    if t: base += f"<{t.group(1)},{t.group(2)},pre{t.group(3)}>"
    return base
t0 = step[0]["ts"]; t1 = max(e["ts"] + e["dur"] for e in step)
agg = collections.defaultdict(lambda: [0, 0.0])
for e in step:
    a = agg[short(e["name"])]; a[0] += 1; a[1] += e["dur"]
print(f"{len(step)} events, span {t1 - t0:.1f} us, sum of durations {sum(e['dur'] for e in step):.1f} us")
for k, (c, d) in sorted(agg.items(), key=lambda kv: -kv[1][1]): print(f"{d:8.1f} n={c:3d} avg {d / c:6.2f} {k}")
if len(sys.argv) > 3:
    for e in step: print(f"{e['ts'] - t0:9.1f} {e['dur']:7.2f} s{e['args']['stream']} {short(e['name'])}")
