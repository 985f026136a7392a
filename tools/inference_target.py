"""ncu target: rollout inference calls (ppo_atari_oc.py:505-506) at `n_envs` Pong-shaped observations through
PPObj.get_action_and_value and nothing else; per call: absmax, thin_fwd_pair_kernel, four tc_gemm_kernel<FWD>, policy_head_kernel.
python tools/inference_target.py [n_envs=4096] [calls=4]"""
import os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oc_cleanrl_b200.architectures.ppo import PPObj
from oc_cleanrl_b200.synthetic import GAME_SHAPES, stub_envs

n = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
calls = int(sys.argv[2]) if len(sys.argv) > 2 else 4
dev = torch.device("cuda:0")
D, A = GAME_SHAPES["pong"]
torch.manual_seed(1)
agent = PPObj(stub_envs(4, D, A), dev, (256, 512, 1024, 512), (512,)).to(dev)
obs = torch.randint(0, 200, (n, 4, D), device=dev).float()
with torch.no_grad():
    for _ in range(calls):
        out = agent.get_action_and_value(obs)
torch.cuda.synchronize()
print("ok", float(out[3].sum()))
