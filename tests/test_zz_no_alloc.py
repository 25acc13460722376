"""GPU: the allocation-free entry point cf_ppo_pi_grad_static_dev against cf_ppo_pi_grad_dev, both through the C ABI.

No counterpart in the reference: the entry point was added for callers that cannot
call cudaMalloc.  Its scratch (the per-CTA partial sums) is a file-scope __device__ array that every call shares and that
keeps the previous call's contents, so the check is made on TWO calls in a row with different inputs and sizes - a large
batch first (all 148 records written), then a small one (one record) whose result must not see the first call's leftovers.
The two entry points run 148 and 296 persistent CTAs, i.e. sum in a different order: 2e-5 of each tensor's largest entry,
the tolerance of tests/test_ppo_grad_kernel.py.  This file is named to be collected last."""
import ctypes as C

import pytest

pytestmark = pytest.mark.gpu

ENV = 'DroneHoverBulletFreeEnvWithRandomHJAdversary-v0'


def _grads(lins):
    import torch
    gw = [torch.full_like(l.weight, 7.0) for l in lins]
    gb = [torch.full_like(l.bias, 7.0) for l in lins]
    return gw, gb


def _ptrs(ts):
    return (C.c_void_p * 3)(*[t.data_ptr() for t in ts])


def test_static_scratch_entry_matches_the_workspace_entry_on_two_calls_in_a_row():
    import torch
    from disturbance_crazyfile_simulation_b200 import abi
    from disturbance_crazyfile_simulation_b200.policy import FusedPolicy, _linears
    from disturbance_crazyfile_simulation_b200.vec_env import make
    from test_policy_kernel import _random_ac
    from test_ppo_grad_kernel import _batch
    L = abi.lib()
    env = make(ENV, num_envs=128, seed=1)
    od, dev = env.obs_dim, env.device
    ac = _random_ac(od, 5, dev)
    fp = FusedPolicy(env, ac)                               # builds the cf_policy_weights struct of device pointers
    lins = _linears(ac.pi.net)
    stream = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    ws = torch.empty(L.cf_ppo_workspace_floats(), dtype=torch.float32, device=dev)
    clip = 0.2
    # everything the two calls need exists before the first one: the entry point under test must not allocate
    cases = []
    for B, seed in ((128 * 148 * 3 + 41, 11), (77, 12)):
        obs, act, adv, logp_old = _batch(ac, od, B, seed, dev, 0.15)
        ref_w, ref_b = _grads(lins)
        new_w, new_b = _grads(lins)
        cases.append((B, obs, act, adv, logp_old, ref_w, ref_b, torch.zeros(2, device=dev), new_w, new_b, torch.zeros(2, device=dev)))
    torch.cuda.synchronize()
    for call, (B, obs, act, adv, logp_old, ref_w, ref_b, ref_lr, new_w, new_b, new_lr) in enumerate(cases):
        args = (C.byref(fp._w), od, C.c_void_p(obs.data_ptr()), C.c_void_p(act.data_ptr()), C.c_void_p(adv.data_ptr()),
                C.c_void_p(logp_old.data_ptr()), B, clip)
        abi.check(L.cf_ppo_pi_grad_dev(*args, _ptrs(ref_w), _ptrs(ref_b), C.c_void_p(ref_lr.data_ptr()),
                                       C.c_void_p(ws.data_ptr()), stream))
        torch.cuda.synchronize()
        free_before = torch.cuda.mem_get_info(dev)[0]
        abi.check(L.cf_ppo_pi_grad_static_dev(*args, _ptrs(new_w), _ptrs(new_b), C.c_void_p(new_lr.data_ptr()), stream))
        torch.cuda.synchronize()
        if call == 1:                                       # the first call per process may load the module lazily
            assert torch.cuda.mem_get_info(dev)[0] == free_before
        assert torch.isfinite(new_lr).all()
        assert abs(float(new_lr[0]) - float(ref_lr[0])) < 2e-5 * max(1.0, abs(float(ref_lr[0]))), call
        assert abs(float(new_lr[1]) - float(ref_lr[1])) < 2e-5, call
        for name, a, b in [(f'w{i}', new_w[i], ref_w[i]) for i in range(3)] + [(f'b{i}', new_b[i], ref_b[i]) for i in range(3)]:
            assert not (a == 7.0).any(), (call, name)       # every entry was written
            d = (a - b).abs().max().item()
            assert d < 2e-5 * max(b.abs().max().item(), 1e-6), (call, name, d, b.abs().max().item())
    # the same small call once more: bit-identical to itself (fixed reduction order, no dependence on the array's old contents)
    B, obs, act, adv, logp_old, _, _, _, new_w, new_b, new_lr = cases[1]
    again_w, again_b = _grads(lins)
    again_lr = torch.zeros(2, device=dev)
    abi.check(L.cf_ppo_pi_grad_static_dev(C.byref(fp._w), od, C.c_void_p(obs.data_ptr()), C.c_void_p(act.data_ptr()),
                                          C.c_void_p(adv.data_ptr()), C.c_void_p(logp_old.data_ptr()), B, clip, _ptrs(again_w),
                                          _ptrs(again_b), C.c_void_p(again_lr.data_ptr()), stream))
    torch.cuda.synchronize()
    for a, b in zip(again_w + again_b + [again_lr], new_w + new_b + [new_lr]):
        assert torch.equal(a, b)
    env.close()


def test_static_scratch_entry_rejects_sizes_beyond_its_limit():
    import torch
    from disturbance_crazyfile_simulation_b200 import abi
    from disturbance_crazyfile_simulation_b200.policy import FusedPolicy, _linears
    from disturbance_crazyfile_simulation_b200.vec_env import make
    from test_policy_kernel import _random_ac
    L = abi.lib()
    limit = L.cf_ppo_static_max_samples()
    assert limit == 1 << 22
    env = make(ENV, num_envs=128, seed=1)
    od, dev = env.obs_dim, env.device
    ac = _random_ac(od, 5, dev)
    fp = FusedPolicy(env, ac)
    lins = _linears(ac.pi.net)
    gw, gb = _grads(lins)
    lr = torch.zeros(2, device=dev)
    small = torch.zeros(128 * od, dtype=torch.float32, device=dev)   # never read: the size check comes before the launch
    rc = L.cf_ppo_pi_grad_static_dev(C.byref(fp._w), od, C.c_void_p(small.data_ptr()), C.c_void_p(small.data_ptr()),
                                     C.c_void_p(small.data_ptr()), C.c_void_p(small.data_ptr()), limit + 1, 0.2, _ptrs(gw), _ptrs(gb),
                                     C.c_void_p(lr.data_ptr()), C.c_void_p(torch.cuda.current_stream(dev).cuda_stream))
    assert rc == abi.CF_ERR_INVALID
    msg = L.cf_last_error().decode()
    assert str(limit + 1) in msg and '4194304' in msg and 'limit' in msg, msg
    torch.cuda.synchronize()
    assert all((g == 7.0).all() for g in gw + gb)           # nothing was launched
    env.close()
