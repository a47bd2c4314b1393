"""Per-tensor error of the device train step against torch fp32/fp64 CPU on the golden batches."""
import os, sys
import numpy as np, torch, torch.nn.functional as F
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from helpers import load_golden_weights
from alphazero_gomoku_b200.engine import Engine, weight_shapes

g = np.load(os.path.join(ROOT, "tests/golden/train_step.npz"))
tag, w, nr = sys.argv[1] if len(sys.argv) > 1 else "8_8_5", 8, 5
if tag == "6_6_4": w, nr = 6, 4
arrs = load_golden_weights("weights_" + tag)
shapes = weight_shapes(w, w)

def fwd(p, x, pi, z, dt):
    x = torch.tensor(x, dtype=dt)
    x = F.relu(F.conv2d(x, p[0], p[1], padding=1)); x = F.relu(F.conv2d(x, p[2], p[3], padding=1)); x = F.relu(F.conv2d(x, p[4], p[5], padding=1))
    a = F.relu(F.conv2d(x, p[6], p[7])).flatten(1); logp = F.log_softmax(F.linear(a, p[8], p[9]), dim=1)
    v = F.relu(F.conv2d(x, p[10], p[11])).flatten(1); v = torch.tanh(F.linear(F.relu(F.linear(v, p[12], p[13])), p[14], p[15]))
    return F.mse_loss(v.view(-1), torch.tensor(z, dtype=dt)) - torch.mean(torch.sum(torch.tensor(pi, dtype=dt) * logp, 1))

e = Engine(width=w, height=w, n_in_row=nr, n_games=1, n_playout=1, nodes_per_game=4)
e.set_weights(arrs)
ps = {dt: [torch.tensor(a, dtype=dt, requires_grad=True) for a in arrs] for dt in (torch.float32, torch.float64)}
opts = {dt: torch.optim.Adam(ps[dt], weight_decay=1e-4) for dt in ps}
for i, lr in enumerate(g["t%s_lrs" % tag]):
    st, pi, z = g["t%s_states" % tag][i].astype(np.float32), g["t%s_pis" % tag][i], g["t%s_zs" % tag][i]
    cur = e.get_weights()
    e.train_backward(st, pi, z)
    gd = e.train_grads()
    q64 = [torch.tensor(a, dtype=torch.float64, requires_grad=True) for a in cur]
    fwd(q64, st, pi, z, torch.float64).backward()
    g_at_dev = np.concatenate([q.grad.numpy().ravel() for q in q64])
    e.train_apply(float(lr))
    grads = {}
    for dt in ps:
        for pg in opts[dt].param_groups: pg["lr"] = float(lr)
        opts[dt].zero_grad()
        fwd(ps[dt], st, pi, z, dt).backward()
        grads[dt] = np.concatenate([q.grad.numpy().ravel().astype(np.float64) for q in ps[dt]])
        opts[dt].step()
    dev = np.concatenate([a.ravel() for a in e.get_weights()]).astype(np.float64)
    p32 = np.concatenate([q.detach().numpy().ravel() for q in ps[torch.float32]]).astype(np.float64)
    p64 = np.concatenate([q.detach().numpy().ravel() for q in ps[torch.float64]])
    print("step", i)
    off = 0
    for k, shp in enumerate(shapes):
        n = int(np.prod(shp)); sl = slice(off, off + n); off += n
        ge_d = np.abs(gd[sl] - g_at_dev[sl]).max(); ge_t = np.abs(grads[torch.float32][sl] - grads[torch.float64][sl]).max()
        gm = np.abs(grads[torch.float64][sl]).max()
        print("  t%02d %-18s |g|max %.2e  grad err dev %.2e torch32 %.2e | param err vs f64: dev %.2e torch32 %.2e | dev vs torch32 %.2e"
              % (k, shp, gm, ge_d, ge_t, np.abs(dev[sl] - p64[sl]).max(), np.abs(p32[sl] - p64[sl]).max(), np.abs(dev[sl] - p32[sl]).max()))
