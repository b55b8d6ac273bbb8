import torch, sys
sys.path.insert(0, "/root/repo")
from tests.test_kuka2_parity_gpu import *
c = k2.Kuka2Consts(); n = 8192; seed = 40 + n
st = k2.make_state(n, seed, c, fresh=False)
task = make_task(n, c); load_state(task, st)
act, rnd = k2.make_actions(n, seed, c), k2.make_rnd(n, seed, c)
before = st["rewards_episode"].clone()
k2.pre_step(st, c, act, rnd); task.ext_rand = rnd.cuda(); task.pre_physics_step(act.cuda())
mid = st["rewards_episode"].clone()
k2.post_step(st, c); task.post_physics_step()
got = dump_state(task)
g, w = got["rewards_episode"], st["rewards_episode"]
bad = (~torch.isclose(g, w, rtol=RTOL, atol=ATOL)).nonzero()
for t, e in bad.tolist():
    print("term", t, "env", e, "got", g[t, e].item(), "want", w[t, e].item(), "before", before[t, e].item(), "mid", mid[t, e].item(),
          "reset", st["reset_buf"][e].item(), "lifted", st["lifted_object"][e].item())
    print(" all terms got", g[:, e].tolist()); print(" all terms want", w[:, e].tolist())
for k in FLOAT:
    gg, ww = got[k].float().reshape(-1), st[k].float().reshape(-1)
    ne = (gg != ww)
    print(k, "bitwise mismatches", int(ne.sum()), "of", gg.numel(), "max", (gg - ww).abs().max().item())
e = 7253
print("ckmd got", got["closest_keypoint_max_dist"][e].item(), "want", st["closest_keypoint_max_dist"][e].item())
o = got["obs_buf"][e]; w = st["obs_buf"][e]
d = (o != w).nonzero().flatten().tolist(); print("obs mismatching cols", d, [(o[i].item(), w[i].item()) for i in d])
