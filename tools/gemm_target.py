"""ncu target: full-size minibatches (131,072 samples of a 4096 x 128 Pong-shaped rollout, default widths, default engine)
through ocppo_minibatch_fwd_bwd and nothing else.  Per minibatch the tensor-core launches are, in order,
FWD l1 l2 l3 l4, then WGRAD l4, DGRAD l4, WGRAD l3, DGRAD l3, WGRAD l2, DGRAD l2, WGRAD l1, DGRAD l1  (l = 0-based trunk
layer; l0 is the CUDA-core first layer), so `ncu -k regex:tc_gemm -s 12 -c 12` captures the second minibatch's twelve.
python tools/gemm_target.py [minibatches=2] [game=pong]"""
import ctypes as C, os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oc_cleanrl_b200 import _abi
from oc_cleanrl_b200.architectures.ppo import PPObj
from oc_cleanrl_b200.learner import LearnerArgs, PPOLearner
from oc_cleanrl_b200.synthetic import GAME_SHAPES, make_rollout, stub_envs

reps = int(sys.argv[1]) if len(sys.argv) > 1 else 2
game = sys.argv[2] if len(sys.argv) > 2 else "pong"
dev = torch.device("cuda:0")
lib = _abi.lib()
F, T, N, mb = 4, 128, 4096, 131072
D, A = GAME_SHAPES[game]
enc, dec = (256, 512, 1024, 512), (512,)
n = T * N
data = make_rollout(T, N, F, D)
rs = np.random.RandomState(1)
flat = (data["obs"].reshape(n, -1).to(dev), torch.as_tensor(rs.randint(0, A, n)).to(dev),
        torch.as_tensor(-np.log(A) + 0.05 * rs.randn(n), dtype=torch.float32).to(dev),
        torch.as_tensor(rs.randn(n), dtype=torch.float32).to(dev), torch.as_tensor(rs.randn(n), dtype=torch.float32).to(dev),
        torch.as_tensor(rs.randn(n), dtype=torch.float32).to(dev))
perm = torch.as_tensor(np.random.RandomState(42).permutation(n), dtype=torch.int64).to(dev)
torch.manual_seed(42)
agent = PPObj(stub_envs(F, D, A), dev, enc, dec).to(dev)
learner = PPOLearner(agent, LearnerArgs(num_envs=N, num_steps=T))
agent._ensure_workspace(mb, training=True, exact=True)
for i in range(reps):
    _abi.check(lib.ocppo_minibatch_fwd_bwd(agent._handle.raw, *[_abi.ptr(t) for t in flat], _abi.ptr(perm[i * mb:(i + 1) * mb]), mb,
                                           C.byref(learner._hp), _abi.ptr(learner._stats[i]), agent._stream()))
torch.cuda.synchronize()
print("loss", [float(x) for x in learner._stats[:reps, 3].cpu()])
