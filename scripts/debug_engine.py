"""GPU debug: compare the native engine's per-layer gradients with torch autograd for one step."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from elephas_b200.models import SGD, zoo

def run(hidden, depth, B, graphs):
    os.environ['ELEPHAS_B200_GRAPHS'] = '1' if graphs else '0'
    torch.manual_seed(0)
    rng = np.random.default_rng(0)
    m = zoo.wide_mlp(input_dim=hidden, hidden=hidden, depth=depth, out_dim=1)
    m.compile(SGD(lr=0.01), 'mse', ['mae'])
    t = zoo.wide_mlp(input_dim=hidden, hidden=hidden, depth=depth, out_dim=1)
    t.compile(SGD(lr=0.01), 'mse', ['mae'])
    t.set_weights(m.get_weights())
    m.to('cuda'); t.to('cuda')
    t._native_engine = lambda *a, **k: None
    x = rng.normal(size=(B, hidden)).astype(np.float32); y = rng.normal(size=(B, 1)).astype(np.float32)
    eng = m._native_engine()
    # one manual step on the engine without the optimizer, to look at gradients
    buf = eng._buf(B)
    eng._to_device_xy(x, y, buf, B)
    eng.stats.zero_()
    eng._forward(buf, training=True); eng._loss(buf, with_grad=True); eng._backward(buf)
    torch.cuda.synchronize()
    xt, yt = torch.from_numpy(x).cuda(), torch.from_numpy(y).cuda()
    params = t.trainable_weights
    out = t(xt, training=True)
    loss = t._loss_fn(yt, out).mean(); loss.backward()
    print("hidden=%d depth=%d B=%d graphs=%d  loss native=%.5f torch=%.5f" % (hidden, depth, B, graphs, eng.stats[0].item()/B, loss.item()))
    for seg, p in zip(eng.layout.segments, params):
        g = eng.layout.view(eng.g32, seg).reshape(seg.shape)
        ref = p.grad
        err = (g - ref).abs().max().item(); mag = ref.abs().max().item()
        print("   %-28s shape=%-14s max|ref|=%.4e max_err=%.4e %s" % (seg.layer.name + ('/k' if seg.index == 0 else '/b'), tuple(seg.shape), mag, err, 'BAD' if err > 0.05 * mag + 1e-4 else ''))
    eng.g32.zero_()
    la = [m.train_on_batch(x, y)[0] for _ in range(8)]
    lb = [t.train_on_batch(x, y)[0] for _ in range(8)]
    print("   native:", np.round(la, 4)); print("   torch :", np.round(lb, 4))

for cfg in [(512, 3, 256, False), (512, 3, 256, True), (256, 4, 256, True), (128, 2, 64, True), (1024, 2, 512, True)]:
    run(*cfg)
