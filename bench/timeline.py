"""In-situ kernel timeline of one replayed training-step graph (CUPTI through torch.profiler).
Writes gpurun_out/timeline.csv: name, stream, start_us, dur_us.  For analysis only - numbers taken
under a profiler are never reported as benchmark results."""
import os, sys, csv
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from distributed_resnet_tensorflow_b200.models.init import init_variables
from distributed_resnet_tensorflow_b200.models.spec import build_spec
from distributed_resnet_tensorflow_b200.train.native import NativeTrainer

def main(batch=128, out="gpurun_out/timeline.csv", dataset="imagenet"):
    spec = build_spec(dataset, 50)
    tr = NativeTrainer(spec, batch, device="cuda:0", weight_decay=1e-4, use_graph=True)
    tr.import_variables(init_variables(spec, seed=0))
    rng = np.random.default_rng(0)
    nbytes = 224 * 224 * 3 if dataset == "imagenet" else 3072
    imgs = torch.from_numpy(rng.integers(0, 256, size=(batch, nbytes), dtype=np.uint8)).pin_memory()
    labs = torch.from_numpy(rng.integers(0, spec.num_classes, size=(batch,), dtype=np.int32)).pin_memory()
    for _ in range(5):
        tr.train_step(imgs, labs, lr=0.1)
    torch.cuda.synchronize()
    from torch.profiler import profile, ProfilerActivity
    with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
        for _ in range(3):
            tr.train_step(imgs, labs, lr=0.1)
        torch.cuda.synchronize()
    rows = []
    for ev in prof.events():
        if ev.device_type == torch.autograd.DeviceType.CUDA:
            rows.append((ev.name[:70], getattr(ev, "stream", -1) if hasattr(ev, "stream") else -1, ev.time_range.start, ev.time_range.end - ev.time_range.start))
    rows.sort(key=lambda r: r[2])
    os.makedirs(os.path.dirname(out), exist_ok=True)
    with open(out, "w", newline="") as f:
        w = csv.writer(f); w.writerow(["name", "stream", "start_us", "dur_us"]); w.writerows(rows)
    print("events", len(rows))
    try:
        prof.export_chrome_trace(out.replace(".csv", ".json"))
    except Exception as e:
        print("trace export failed", e)

if __name__ == "__main__":
    ds = sys.argv[1] if len(sys.argv) > 1 else "imagenet"
    b = int(sys.argv[2]) if len(sys.argv) > 2 else 128
    main(batch=b, out=f"gpurun_out/timeline_{ds}_{b}.csv", dataset=ds)
