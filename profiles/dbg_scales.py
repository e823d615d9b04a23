import ctypes as C, torch, sys
sys.path.insert(0,'/root/repo')
from legged_tracking_b200 import _abi, ppo
from tests.test_gpu_policy import _random_batch, _batch_ptrs, _p
lib=_abi.load()
torch.manual_seed(11)
dev=torch.device('cuda',0)
ac=ppo.ActorCritic(30,5,90,12).to(dev)
B=400;count=256
b=_random_batch(ac,B,5,dev)
idx=torch.randperm(B)[:count].to(dev)
ac.grads=torch.zeros(ac.num_params,device=dev)
ac._ready(count,training=True)
stats=torch.zeros(16,device=dev)
hp=ppo.PPO.hparams(); bp=_batch_ptrs(b)
for it in range(3):
    _abi.check_policy(lib.lt_ppo_minibatch_grad(ac._h,C.byref(bp),_p(idx),count,C.byref(hp),1,_p(stats),None),ac._h)
    torch.cuda.synchronize()
    ptr=lib.lt_policy_grad_scales(ac._h)
    host=(C.c_float*30)()
    C.CDLL("libcudart.so.12").cudaMemcpy(C.c_void_p(C.addressof(host)),C.c_void_p(ptr),C.c_size_t(120),C.c_int(2))
    print([f"{x:.3g}" for x in host][:15])
    lo,hi=ac.body_slices['adaptation_module']
    print(float(ac.grads[lo:hi].abs().max()), float(ac.grads[hi:].abs().max()))
