"""numpy prototype of the twisted (two-sided) local solve used by the TMEM fused kernel:
lane block of N=31 interior cells; cells 0..M-1 eliminated top-down, cells N-1..M bottom-up."""
import numpy as np, sys
sys.path.insert(0, '/root/repo'); sys.path.insert(0, '/root/repo/scratch')
from oracle import fvm_oracle as O
from fvm_b200 import workloads as W
import proto_warp_spike as P0

L, C, N, M = 32, 32, 31, 16

def setup(alo, adi, aup, theta):
    alo = alo.reshape(L, C); adi = adi.reshape(L, C); aup = aup.reshape(L, C)
    rth, cth = 1/theta, (1-theta)/theta
    e = np.zeros((L, N)); s = np.zeros((L, N)); b = np.zeros((L, N))
    gv = np.zeros((L, N)); gu = np.zeros((L, N))
    sprev = np.zeros(L)
    for c in range(M):                       # left: top-down
        inv = 1.0 / (adi[:, c] - alo[:, c] * sprev)
        e[:, c] = alo[:, c] * inv; s[:, c] = aup[:, c] * inv; b[:, c] = inv * (rth - cth * adi[:, c])
        sprev = s[:, c]
        gv[:, c] = e[:, c] if c == 0 else -e[:, c] * gv[:, c-1]
    sprev = np.zeros(L)
    for c in range(N-1, M-1, -1):            # right: bottom-up
        inv = 1.0 / (adi[:, c] - aup[:, c] * sprev)
        e[:, c] = aup[:, c] * inv; s[:, c] = alo[:, c] * inv; b[:, c] = inv * (rth - cth * adi[:, c])
        sprev = s[:, c]
        gu[:, c] = e[:, c] if c == N-1 else -e[:, c] * gu[:, c+1]
    delta = 1.0 / (1.0 - s[:, M] * s[:, M-1])
    def subst(g):
        x = g.copy()
        x[:, M] = (x[:, M] - s[:, M] * x[:, M-1]) * delta
        for c in range(M-1, -1, -1): x[:, c] = x[:, c] - s[:, c] * x[:, c+1]
        for c in range(M+1, N): x[:, c] = x[:, c] - s[:, c] * x[:, c-1]
        return x
    v = subst(gv); u = subst(gu)
    slo = alo[:, C-1]; sdi = adi[:, C-1]; sup = aup[:, C-1]
    nxt = lambda x: np.concatenate([x[1:], [0.0]])
    RL = -slo * v[:, N-1]; RD = sdi - slo * u[:, N-1] - sup * nxt(v[:, 0]); RU = -sup * nxt(u[:, 0])
    al = []; ga = []; Lc, Dc, Uc = RL.copy(), RD.copy(), RU.copy(); st = 1
    def sh(x, k, fill):
        out = np.full(L, fill, dtype=float)
        if k > 0: out[k:] = x[:-k]
        else: out[:k] = x[-k:]
        return out
    while st < L:
        a_ = -Lc / sh(Dc, st, 1.0); g_ = -Uc / sh(Dc, -st, 1.0)
        al.append(a_); ga.append(g_)
        Dn = Dc + a_ * sh(Uc, st, 0.0) + g_ * sh(Lc, -st, 0.0)
        Lc, Uc = a_ * sh(Lc, st, 0.0), g_ * sh(Uc, -st, 0.0); Dc = Dn; st *= 2
    return dict(e=e, s=s, b=b, v=v, u=u, delta=delta, slo=slo, sup=sup, sblo=-cth*slo, sbdi=rth-cth*sdi,
                sbup=-cth*sup, al=al, ga=ga, invD=1.0/Dc, cth=cth)

def step(cf, w):
    w = w.reshape(L, C).copy()
    wl = np.concatenate([[0.0], w[:-1, C-1]]); wr = np.concatenate([w[1:, 0], [0.0]])
    old = w.copy()
    e, s, b, cth = cf['e'], cf['s'], cf['b'], cf['cth']
    g = np.zeros((L, N))
    for c in range(M):
        wm = wl if c == 0 else old[:, c-1]
        t = s[:, c] * old[:, c+1] + e[:, c] * wm
        r = b[:, c] * old[:, c] - cth * t
        g[:, c] = r if c == 0 else r - e[:, c] * g[:, c-1]
    for c in range(N-1, M-1, -1):
        t = s[:, c] * old[:, c-1] + e[:, c] * old[:, c+1]
        r = b[:, c] * old[:, c] - cth * t
        g[:, c] = r if c == N-1 else r - e[:, c] * g[:, c+1]
    rs = cf['sblo'] * old[:, N-1] + cf['sbdi'] * old[:, C-1] + cf['sbup'] * wr
    g[:, M] = (g[:, M] - s[:, M] * g[:, M-1]) * cf['delta']
    for c in range(M-1, -1, -1): g[:, c] -= s[:, c] * g[:, c+1]
    for c in range(M+1, N): g[:, c] -= s[:, c] * g[:, c-1]
    rr = rs - cf['slo'] * g[:, N-1] - cf['sup'] * np.concatenate([g[1:, 0], [0.0]])
    st = 1
    for a_, g_ in zip(cf['al'], cf['ga']):
        rm = np.zeros(L); rm[st:] = rr[:-st]; rp = np.zeros(L); rp[:-st] = rr[st:]
        rr = rr + a_ * rm + g_ * rp; st *= 2
    xs = rr * cf['invD']; xsl = np.concatenate([[0.0], xs[:-1]])
    x = np.zeros((L, C))
    x[:, :N] = g - cf['v'] * xsl[:, None] - cf['u'] * xs[:, None]
    x[:, C-1] = xs
    return x.reshape(-1)

if __name__ == '__main__':
    wl = W.batched_sweep(65536, 1024, count=3)
    for p in range(3):
        asm = O.assemble(wl['faces'], wl['a'][p], wl['d'][p], 'exponential', **wl['bc'])
        A, B = O.theta_operators(asm, wl['k'][p], 0.5)
        cf = setup(A['lo'], A['di'], A['up'], 0.5)
        ref = O.theta_run_scipy(asm, wl['k'][p], 0.5, wl['w0'][p], 200)
        w = wl['w0'][p].copy(); err = 0
        for n in range(200):
            w = step(cf, w); err = max(err, np.max(np.abs(w - ref[n])) / np.max(np.abs(ref[n])))
        print('cfg3 p', p, err)
    # ragged J with Dirichlet
    J = 777; f = W.graded_faces(J, 10.0); c = 0.5*(f[1:]+f[:-1]); w0 = W.pulse(c) + 0.2
    asm = O.assemble(f, -1.3, 0.004, 'upwind', left_value=1., right_value=0.)
    for theta in (0.5, 1.0, 0.7):
        A, B = O.theta_operators(asm, 2e-3, theta)
        pad = lambda x, fill: np.concatenate([x, np.full(1024 - J, fill)])
        cf = setup(pad(A['lo'], 0.), pad(A['di'], 1.), pad(A['up'], 0.), theta)
        ref = O.theta_run_scipy(asm, 2e-3, theta, w0, 60)
        w = pad(w0, 0.0); err = 0
        for n in range(60):
            w = step(cf, w); err = max(err, np.max(np.abs(w[:J] - ref[n])) / np.max(np.abs(ref[n])))
        print('ragged theta', theta, err)
