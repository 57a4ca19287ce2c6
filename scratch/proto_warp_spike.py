"""numpy prototype of the fused kernel's per-warp algorithm (32 lanes x C cells,
last cell of each lane = separator; precomputed local LU + spikes + PCR on the
32 separators; 5 coefficients per interior cell)."""
import numpy as np, sys
sys.path.insert(0, '/root/repo')
from oracle import fvm_oracle as O
from fvm_b200 import workloads as W

def setup(alo, adi, aup, theta, C):
    """alo/adi/aup: (32*C,) diagonals of A (padded rows identity). returns coef dict."""
    L = 32
    alo = alo.reshape(L, C); adi = adi.reshape(L, C); aup = aup.reshape(L, C)
    n = C - 1
    p = np.zeros((L, n)); q = np.zeros((L, n)); inv = np.zeros((L, n))
    for c in range(n):
        den = adi[:, c] - (alo[:, c] * q[:, c-1] if c > 0 else 0.0)
        inv[:, c] = 1.0 / den
        p[:, c] = alo[:, c] * inv[:, c]
        q[:, c] = aup[:, c] * inv[:, c]
    # spikes
    v = np.zeros((L, n)); u = np.zeros((L, n))
    if n > 0:
        gv = np.zeros((L, n)); gv[:, 0] = alo[:, 0] * inv[:, 0]
        for c in range(1, n): gv[:, c] = -p[:, c] * gv[:, c-1]
        v[:, n-1] = gv[:, n-1]
        for c in range(n-2, -1, -1): v[:, c] = gv[:, c] - q[:, c] * v[:, c+1]
        u[:, n-1] = q[:, n-1]
        for c in range(n-2, -1, -1): u[:, c] = -q[:, c] * u[:, c+1]
    cth = (1 - theta) / theta
    bdi = inv * (1 / theta - cth * adi[:, :n])
    # separator rows
    slo = alo[:, C-1]; sdi = adi[:, C-1]; sup = aup[:, C-1]
    nxt = lambda x: np.concatenate([x[1:], [0.0]])
    if n > 0:
        RL = -slo * v[:, n-1]
        RD = sdi - slo * u[:, n-1] - sup * nxt(v[:, 0])
        RU = -sup * nxt(u[:, 0])
    else:
        RL, RD, RU = slo.copy(), sdi.copy(), sup.copy()
    # PCR factorisation
    al = []; ga = []
    Lc, Dc, Uc = RL.copy(), RD.copy(), RU.copy()
    s = 1
    while s < L:
        def sh(x, k, fill):
            out = np.full(L, fill, dtype=float)
            if k > 0: out[k:] = x[:-k]
            else: out[:k] = x[-k:]
            return out
        Dm = sh(Dc, s, 1.0); Dp = sh(Dc, -s, 1.0)
        Lm = sh(Lc, s, 0.0); Um = sh(Uc, s, 0.0); Lp = sh(Lc, -s, 0.0); Up = sh(Uc, -s, 0.0)
        a_ = -Lc / Dm; g_ = -Uc / Dp
        al.append(a_); ga.append(g_)
        Dc = Dc + a_ * Um + g_ * Lp
        Lc = a_ * Lm; Uc = g_ * Up
        s *= 2
    return dict(p=p, q=q, bdi=bdi, v=v, u=u, slo=slo, sup=sup,
                sblo=-cth*slo, sbdi=1/theta - cth*sdi, sbup=-cth*sup,
                al=al, ga=ga, invD=1.0/Dc, cth=cth, C=C)

def step(cf, w):
    L = 32; C = cf['C']; n = C - 1
    w = w.reshape(L, C)
    wl = np.concatenate([[0.0], w[:-1, C-1]])   # left neighbour of c=0: prev lane's separator
    wr = np.concatenate([w[1:, 0], [0.0]])       # right neighbour of separator: next lane's c=0
    g = np.zeros((L, n))
    for c in range(n):
        wm = wl if c == 0 else w[:, c-1]
        wp = w[:, c+1]
        t = cf['p'][:, c] * wm + cf['q'][:, c] * wp
        rt = cf['bdi'][:, c] * w[:, c] - cf['cth'] * t
        g[:, c] = rt - (cf['p'][:, c] * g[:, c-1] if c > 0 else 0.0)
    for c in range(n-2, -1, -1):
        g[:, c] = g[:, c] - cf['q'][:, c] * g[:, c+1]
    wsm = w[:, C-2] if n > 0 else wl
    rs = cf['sblo'] * wsm + cf['sbdi'] * w[:, C-1] + cf['sbup'] * wr
    if n > 0:
        g0n = np.concatenate([g[1:, 0], [0.0]])
        rr = rs - cf['slo'] * g[:, n-1] - cf['sup'] * g0n
    else:
        rr = rs
    s = 1
    for a_, g_ in zip(cf['al'], cf['ga']):
        rm = np.zeros(L); rm[s:] = rr[:-s]
        rp = np.zeros(L); rp[:-s] = rr[s:]
        rr = rr + a_ * rm + g_ * rp
        s *= 2
    xs = rr * cf['invD']
    xsl = np.concatenate([[0.0], xs[:-1]])
    x = np.zeros((L, C))
    for c in range(n):
        x[:, c] = g[:, c] - cf['v'][:, c] * xsl - cf['u'][:, c] * xs
    x[:, C-1] = xs
    return x.reshape(-1)

def run_case(faces, a, d, k, scheme, bc, theta, w0, nsteps, C):
    asm = O.assemble(faces, a, d, scheme, **bc)
    A, B = O.theta_operators(asm, k, theta)
    J = len(w0); Jp = 32 * C
    pad = lambda x, fill: np.concatenate([x, np.full(Jp - J, fill)])
    cf = setup(pad(A['lo'], 0.), pad(A['di'], 1.), pad(A['up'], 0.), theta, C)
    w = pad(w0, 0.0)
    ref = O.theta_run_scipy(asm, k, theta, w0, nsteps)
    err = 0
    for n in range(nsteps):
        w = step(cf, w)
        e = np.max(np.abs(w[:J] - ref[n])) / max(np.max(np.abs(ref[n])), 1e-6*np.max(np.abs(w0)))
        err = max(err, e)
    return err

if __name__ == '__main__':
    wl = W.batched_sweep(65536, 1024, count=4)
    for p in range(4):
        print('cfg3 p', p, run_case(wl['faces'], wl['a'][p], wl['d'][p], wl['k'][p], 'exponential', wl['bc'], 0.5, wl['w0'][p], 300, 32))
    f = O.example_faces('geometric'); c = 0.5*(f[1:]+f[:-1]); w0 = np.sin(np.pi*c)**100
    for scheme in O.SCHEMES:
        for theta in (0.5, 1.0):
            print('cfg1', scheme, theta, run_case(f, 1, 1e-3, 0.01, scheme, dict(left_value=1., right_value=0.), theta, w0, 251, 2))
    f = O.demo_faces(); c = 0.5*(f[1:]+f[:-1]); w0 = np.sin(np.pi*c)**100; w0[0] = 1
    print('demo', run_case(f, 1, 1e-3, 0.01, 'exponential', dict(left_value=1., right_value=0.), 1.0, w0, 201, 8))
    f = np.linspace(0,1,33); c = 0.5*(f[1:]+f[:-1]); w0 = np.sin(np.pi*c)**2
    print('C=1', run_case(f, 1, 1e-1, 0.01, 'upwind', dict(left_flux=0., right_flux=0.), 0.5, w0, 50, 1))
