"""Each GAE shape in its own process (a trapped kernel kills the CUDA context): python tools/gae_cases.py"""
import subprocess, sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CASES = [(128, 4096), (5, 4), (1, 4), (31, 32), (32, 32), (33, 36), (128, 8), (129, 36), (257, 100), (64, 160), (700, 44), (300, 64), (5, 64), (40, 4)]
if len(sys.argv) == 3:
    import torch
    sys.path.insert(0, ROOT)
    from oc_cleanrl_b200.learner import compute_gae
    from oc_cleanrl_b200.synthetic import make_gae_inputs
    from oracle import ref_ppo
    T, N = int(sys.argv[1]), int(sys.argv[2])
    d = make_gae_inputs(T, N, seed=42)
    a_ref, r_ref = ref_ppo.ref_gae(d["rewards"], d["values"], d["dones"], d["next_value"], d["next_done"], 0.99, 0.95)
    g = {k: v.cuda() for k, v in d.items()}
    adv, ret = compute_gae(g["rewards"], g["values"], g["dones"], g["next_value"], g["next_done"], 0.99, 0.95)
    torch.cuda.synchronize()
    print("exact" if torch.equal(adv.cpu(), a_ref) and torch.equal(ret.cpu(), r_ref) else f"MISMATCH {(adv.cpu() - a_ref).abs().max()}")
else:
    for T, N in CASES:
        r = subprocess.run([sys.executable, __file__, str(T), str(N)], capture_output=True, text=True, timeout=120)
        tail = (r.stdout.strip().splitlines() or [""])[-1] if r.returncode == 0 else (r.stderr.strip().splitlines() or ["?"])[-1][:150]
        print(f"T={T} N={N}: rc={r.returncode} {tail}", flush=True)
