"""Per-tensor gradient error of ONE full-size minibatch (131,072 samples of a 4096 x 128 rollout, default widths) over
rollout seeds x games x GEMM engines, against the oracle's operator sequence run by torch on the device in fp32 and fp64
(tests/fullsize.py).  Usage: python tools/fullsize_sweep.py [--engines 8,6] [--seeds 5] [--games pong,spaceinvaders,breakout]
Seeds >= 3 run on "trained-like" weights (every tensor disturbed, biases non-zero)."""
import argparse, os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from fullsize import fullsize_case  # noqa: E402

p = argparse.ArgumentParser()
p.add_argument("--engines", default="8,6")
p.add_argument("--seeds", type=int, default=5)
p.add_argument("--games", default="pong,spaceinvaders,breakout")
p.add_argument("--masks", action="store_true", help="also against fp64 under the device path's ReLU decisions")
a = p.parse_args()
dev = torch.device("cuda:0")
fmt = lambda xs: " ".join(f"{x:.1e}" for x in xs)
worst = {}
print("# tensors in parameters() order: network.0.weight, .bias, network.2.*, network.4.*, network.6.*, network.9.*, actor.*, critic.*")
for engine in [int(e) for e in a.engines.split(",")]:
    for game in a.games.split(","):
        for seed in range(a.seeds):
            r = fullsize_case(dev, game, seed, engine, trained_like=seed >= 3, with_masks=a.masks, with_twin=True)
            tag = f"engine {engine} {game} seed {seed}{' trained-like' if seed >= 3 else ''}"
            print(f"{tag}: vs torch fp32 (max-norm)  {fmt(r['e32'])}")
            print(f"{tag}: vs torch fp64 (max-norm)  {fmt(r['e64'])}")
            print(f"{tag}: torch fp32 vs fp64        {fmt(r['ref'])}")
            print(f"{tag}: torch fp32, units permuted, vs torch fp32  {fmt(r['twin32'])}")
            print(f"{tag}: vs torch fp32 (per row)   {fmt(r['rows32'])}")
            print(f"{tag}: torch fp32 vs fp64 (row)  {fmt(r['rows_ref'])}")
            if a.masks:
                print(f"{tag}: vs fp64, device decisions {fmt(r['e64m'])}")
            sys.stdout.flush()
            w = worst.setdefault(engine, dict(e32=0, e64=0, ref=0, rows32=0, rows_ref=0, twin32=0, case32=None))
            w["twin32"] = max(w["twin32"], max(r["twin32"]))
            if max(r["e32"]) > w["e32"]:
                w["e32"], w["case32"] = max(r["e32"]), tag
            for k in ("e64", "ref", "rows32", "rows_ref"):
                w[k] = max(w[k], max(r[k]))
for engine, w in worst.items():
    print(f"WORST engine {engine}: vs torch fp32 {w['e32']:.2e} ({w['case32']}); vs fp64 {w['e64']:.2e}; torch fp32 vs fp64 {w['ref']:.2e}; torch fp32 vs torch fp32 with permuted units {w['twin32']:.2e}; "
          f"per-row vs fp32 {w['rows32']:.2e} (torch fp32 vs fp64 per row: {w['rows_ref']:.2e})")
