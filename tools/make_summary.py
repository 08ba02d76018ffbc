"""Builds profiles/r1_summary.md, the per-run JSON copies and profiles/r1_traffic.json from a set of gpurun outputs:
python tools/make_summary.py <prefix>   (reads gpurun_out/<prefix>_*.json / .csv / .ncu-rep)"""
import io
import json
import os
import shutil
import sys
from contextlib import redirect_stdout

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tools"))
import ncu_table  # noqa: E402

pre = sys.argv[1]
G = os.path.join(ROOT, "gpurun_out")
P = os.path.join(ROOT, "profiles")


def load(name):
    path = os.path.join(G, f"{pre}_{name}.json")
    if not os.path.exists(path):
        return None
    txt = open(path).read().strip().splitlines()
    return json.loads(txt[-1]) if txt else None


def keep(name, dst):
    src = os.path.join(G, f"{pre}_{name}")
    if os.path.exists(src):
        shutil.copy(src, os.path.join(P, dst))


def ktable(d):
    out = ["| kernel | launches/step | us/launch | us/step |", "|---|---|---|---|"]
    for k, v in d.get("kernels", {}).items():
        out.append(f"| `{k}` | {v['launches_per_step']:.0f} | {v['ms_per_launch'] * 1e3:.1f} | {v['ms_per_step'] * 1e3:.1f} |")
    return "\n".join(out)


def table_of(rep, js):
    buf = io.StringIO()
    with redirect_stdout(buf):
        ncu_table.main(os.path.join(G, rep), os.path.join(P, js))
    return buf.getvalue()


c2, r2, c3, r3, c4 = load("bench_c2"), load("ref_c2"), load("bench_c3"), load("ref_c3"), load("bench_c4")
g2 = load("bench_c2_2gpu")
for n, d in (("bench_c2.json", "r1_bench_c2_b200.json"), ("ref_c2.json", "r1_bench_c2_reference.json"),
             ("bench_c3.json", "r1_bench_c3_b200.json"), ("ref_c3.json", "r1_bench_c3_reference.json"),
             ("bench_c4.json", "r1_bench_c4_1gpu.json"), ("bench_c2_2gpu.json", "r1_bench_c2_gnn_2gpu.json"),
             ("launches_c2.csv", "r1_launches_c2.csv")):
    keep(n, d)

L = []
A = L.append
clk = c2.get("clocks", {})
A(f"# Round 1 measurements (B200, 148 SMs, SM clock {clk.get('sm_mhz')} MHz under load, throttle reasons {clk.get('reasons')}; "
  "every number below is from `gpurun` runs of this repo)\n")
A("Raw JSON lines: `profiles/r1_bench_*.json` (bench.py, CUDA events on the launching stream, L2 flushed between steps).")
A("Launch list of the default bench command: `profiles/r1_launches_c2.csv` (`ncu --metrics gpu__time_duration.sum --clock-control none`).")
A("ncu `--set full` tables: `profiles/r1_ncu_*.json` (`tools/ncu_table.py`); the `.ncu-rep` files stay in `gpurun_out/` (scratch).")
A("Boxes differ by a few per cent from run to run (0.362 .. 0.397 ms for the same C2 binary); both arms of a table come from one box.\n")
A("## Headline: rasterizer fwd+bwd, C2 (20x20 cloth, 10 108 Gaussians, SH 3, 4 views 800x800)\n")
A("| arm | device-resident Mpix/s | ms/step | e2e Mpix/s (host buffers) | e2e ms/step |\n|---|---|---|---|---|")
A(f"| this repo | {c2['value']:.0f} | {c2['ms_per_step']:.3f} | {c2['e2e']['value']:.0f} | {c2['e2e']['ms_per_step']:.3f} |")
A(f"| reference CUDA (`oracle/_ref`, same B200) | {r2['value']:.0f} | {r2['ms_per_step']:.3f} | {r2['e2e']['value']:.0f} | {r2['e2e']['ms_per_step']:.3f} |")
A(f"| ratio | {c2['value'] / r2['value']:.2f}x | | {c2['e2e']['value'] / r2['e2e']['value']:.2f}x | |\n")
cb = c2.get("cpu_baseline")
if cb:
    A(f"CPU baseline ({cb['kind']}, {cb['cores']} host threads): {cb['value']:.1f} Mpix/s -- {cb['sample']}.\n")
if g2:
    A(f"2 GPUs (views x2, one flat all-reduce of the gradient bucket): {g2['value']:.0f} Mpix/s, "
      f"{g2['ms_per_step']:.3f} ms/step (e2e {g2['e2e']['value']:.0f} Mpix/s).\n")
A("Per-kernel CUDA-event times inside the step (`kernels{}` of the bench line; event bracketing adds ~3 us to every short kernel, "
  "the launch list has the undisturbed durations):\n")
A(ktable(c2) + "\n")
scal = {}
for n in (1, 2, 4, 8):
    d = load(f"scale_n{n}")
    if d:
        scal[n] = d
if len(scal) > 1 and 1 in scal:
    A("## Weak scaling over views (C2, 4 views per GPU, gradient bucket summed by `csb_peer_all_reduce` over NVLink peer memory)\n")
    A("| GPUs | Mpix/s (device-resident step) | ms/step | efficiency vs N x 1 GPU | e2e Mpix/s | all-reduce kernel: wait for peers / reduce+push / depart (us) | GNN C5 training graphs/s (128 graphs per GPU) |\n|---|---|---|---|---|---|---|")
    for n, d in sorted(scal.items()):
        ar = d.get("allreduce") or {}
        ph = f"{ar.get('wait_peers_us', 0):.1f} / {ar.get('reduce_push_us', 0):.1f} / {ar.get('depart_us', 0):.1f}" if ar.get("calls") else "-"
        tr = (d.get("gnn") or {}).get("train_c5")
        tr1 = (scal[1].get("gnn") or {}).get("train_c5")
        gt = f"{tr['graphs_per_s']:.0f} ({100 * tr['graphs_per_s'] / (n * tr1['graphs_per_s']):.0f} %)" if tr and tr1 else "-"
        A(f"| {n} | {d['value']:.0f} | {d['ms_per_step']:.4f} | {100 * d['value'] / (n * scal[1]['value']):.1f} % | {d['e2e']['value']:.0f} | {ph} | {gt} |")
    sp = scal[1].get("step_ms_spread")
    if sp:
        A(f"\nSpread of the timed steps on one GPU: min {sp['min']:.4f} / median {sp['median']:.4f} / max {sp['max']:.4f} ms, sigma {sp['std'] * 1e3:.1f} us "
          "-- the synchronous all-reduce turns this into the `wait for peers` column (rank skew), which is the larger part of the loss at 8 GPUs.")
    A("All runs of this table come from ONE 8-GPU box (`tools/scale_run_full.sh`: the default `bench.py` command at every N); e2e (host buffers over PCIe) is bounded by the host's memory system when 8 ranks copy at once.\n")
    json.dump({str(n): {k: d.get(k) for k in ("value", "ms_per_step", "n_gpus", "e2e", "allreduce", "step_ms_spread", "config", "clocks")} | {"gnn_train_c5": (d.get("gnn") or {}).get("train_c5")}
               for n, d in scal.items()}, open(os.path.join(P, "r1_scaling_c2.json"), "w"), indent=1)
rf = c2["roofline"]
A(f"`roofline` of the bench line: dominant kernel `{rf['kernel']}`, algorithmic {rf['achieved']:.0f} GB/s of {rf['peak']:.0f} GB/s measured "
  f"HBM peak (frac {rf['frac']:.3f}), ncu DRAM traffic per launch {rf.get('traffic')} B -- the compositing kernels are issue-bound, see the ncu table.\n")
if c3 and r3:
    A("## C3 stress (1 M Gaussians, SH 3, 1920x1080)\n")
    A(f"this repo {c3['value']:.0f} Mpix/s ({c3['ms_per_step']:.2f} ms) vs reference CUDA {r3['value']:.0f} Mpix/s ({r3['ms_per_step']:.2f} ms): "
      f"{c3['value'] / r3['value']:.2f}x; e2e {c3['e2e']['value']:.0f} vs {r3['e2e']['value']:.0f} Mpix/s "
      f"({c3['config'].get('instances_per_step')} instances per step).\n")
    A(ktable(c3) + "\n")
if c4:
    A("## C4 (100 k Gaussians x 64 views 800x800 on one GPU)\n")
    A(f"{c4['value']:.0f} Mpix/s ({c4['ms_per_step']:.1f} ms/step, 4 launch sets of 16 views).\n")
g, rg = c2.get("gnn"), r2.get("gnn")
if g and rg:
    A("## GNN (meshnet, 15 message-passing steps, bf16 tcgen05)\n")
    A("| case | this repo | reference op sequence (fp32 torch on the same B200) | ratio |\n|---|---|---|---|")
    a, b = g["c1_20x20"], rg["c1_20x20"]
    A(f"| C1 forward, 1 graph 20x20 | {a['ms_per_forward']:.3f} ms = {a['graphs_per_s']:.0f} graphs/s (e2e {a['e2e_graphs_per_s']:.0f}) | "
      f"{b['ms_per_forward']:.3f} ms = {b['graphs_per_s']:.0f} graphs/s | {a['graphs_per_s'] / b['graphs_per_s']:.1f}x |")
    a, b = g["batch128_40x40"], rg["batch128_40x40"]
    A(f"| forward, 128 graphs 40x40 (N={a['nodes']}, E={a['edges']}) | {a['ms_per_forward']:.2f} ms = {a['graphs_per_s']:.0f} graphs/s, "
      f"{a['tflops']:.0f} TFLOP/s = {100 * a['roofline']['frac']:.1f} % of measured bf16 peak | {b['ms_per_forward']:.1f} ms = {b['graphs_per_s']:.0f} graphs/s | "
      f"{a['graphs_per_s'] / b['graphs_per_s']:.1f}x |")
    a, b = g.get("train_c5"), rg.get("train_c5")
    if a and b:
        A(f"| C5 share: training step, {a['graphs']} graphs 40x40, 10-step rollout loss, fwd+bwd+Adam | {a['ms_per_step']:.0f} ms = {a['graphs_per_s']:.1f} graphs/s, "
          f"{a['tflops']:.0f} TFLOP/s | {b['graphs_per_s']:.1f} graphs/s on a {b['graphs']}-graph sample (autograd activations of 128 graphs do not fit) | "
          f"{a['graphs_per_s'] / b['graphs_per_s']:.1f}x |")
    if g2 and g2.get("gnn", {}).get("train_c5"):
        t = g2["gnn"]["train_c5"]
        A(f"| same, 2 GPUs (graphs x2, one flat all-reduce of 2.33 M gradients) | {t['graphs_per_s']:.1f} graphs/s = "
          f"{100 * t['graphs_per_s'] / (2 * g['train_c5']['graphs_per_s']):.0f} % efficiency | | |")
    cpu = g.get("cpu_baseline")
    if cpu:
        A(f"| CPU baseline ({cpu['cores']} host threads, fp32 torch port of the reference forward, C1) | | {cpu['value']:.0f} graphs/s | |")
    A("")
t, rt = c2.get("refine_tail"), r2.get("refine_tail")
if t and rt:
    A("## Photometric loss + Adam (the tail of the refinement step, SURVEY 8(f) #1 and #3)\n")
    A("| op | this repo | reference op sequence (torch on the same B200) | ratio |\n|---|---|---|---|")
    a, b = t["image_loss"], rt["image_loss"]
    A(f"| `image_losses` fwd+bwd, 4x3x800x800, lambda 0.1 | {a['ms_fwd_bwd'] * 1e3:.0f} us | {b['ms_fwd_bwd'] * 1e3:.0f} us | {b['ms_fwd_bwd'] / a['ms_fwd_bwd']:.1f}x |")
    a2, b2 = t["adam"], rt["adam"]
    A(f"| Adam step, 7 groups, {a2['elements']} elements | {a2['ms_per_step'] * 1e3:.1f} us | {b2['ms_per_step'] * 1e3:.0f} us | {b2['ms_per_step'] / a2['ms_per_step']:.0f}x |")
    A("")
    for grp in (a, a2):
        for k, v in grp.get("kernels", {}).items():
            r = v["roofline"]
            A(f"* `{k}`: {v['ms_per_launch'] * 1e3:.1f} us/launch, algorithmic {r['achieved']:.0f} GB/s = {100 * r['frac']:.0f} % of the measured HBM peak")
    A("")
if os.path.exists(os.path.join(G, f"{pre}_gnn_edge_full.ncu-rep")):
    A("## ncu `--set full --clock-control none`: GNN edge role, two tiles in flight per CTA (128 graphs of 40x40, one message-passing step)\n")
    A(table_of(f"{pre}_gnn_edge_full.ncu-rep", "r1_ncu_gnn_edge.json"))
    kt = (c2.get("gnn") or {}).get("train_c5", {}).get("kernels")
    if kt:
        A("Kernels of one C5 training step on one GPU (128 graphs, 10-step rollout; CUDA events of the library profiler):\n")
        A("| kernel | launches | ms/step | us/launch |\n|---|---|---|---|")
        for k, v in sorted(kt.items(), key=lambda kv: -kv[1]["ms_per_step"]):
            A(f"| `{k}` | {v['launches']} | {v['ms_per_step']:.1f} | {v['ms_per_step'] / v['launches'] * 1e3:.0f} |")
        A("")
traffic = {}
for title, rep, js, wl in (("C2 step", f"{pre}_c2_full.ncu-rep", "r1_ncu_c2_step.json", "c2"),
                           ("C3 step", f"{pre}_c3_full.ncu-rep", "r1_ncu_c3_step.json", "c3"),
                           ("image loss + Adam", f"{pre}_loss_full.ncu-rep", "r1_ncu_loss_adam.json", None)):
    if os.path.exists(os.path.join(G, rep)):
        A(f"## ncu `--set full --clock-control none`: {title}\n")
        A(table_of(rep, js))
        if wl:
            rows = json.load(open(os.path.join(P, js)))
            tr = {}
            for r in rows:
                name = r["kernel"].replace("_kernel", "").split("<")[0]
                b = (r["rd MB"] + r["wr MB"]) * 1e6
                tr[name] = max(tr.get(name, 0), b)
            traffic[wl] = tr
if traffic:
    json.dump(traffic, open(os.path.join(P, "r1_traffic.json"), "w"), indent=1)
notes = os.path.join(P, "r1_notes.md")
if os.path.exists(notes):
    A(open(notes).read())
open(os.path.join(P, "r1_summary.md"), "w").write("\n".join(L) + "\n")
print("wrote profiles/r1_summary.md")
