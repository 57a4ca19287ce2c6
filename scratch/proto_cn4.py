"""numpy prototype of the 4-operation Crank-Nicolson sweep (theta = 1/2, c = 1):
A^-1 B = (1/theta) A^-1 - c I, so a step on the block interior is  G = (1/theta) T^-1 g - g.
Unnormalised elimination h_c = g_c - l_c h_prev (1 FMA), then with z = y - g:
t_c = h_c*kappa_c - g_c - s_c*g_next (2 FMA) and z_c = t_c - s_c z_next (1 FMA).
On top of proto_pipe's delayed correction: P1 = -(1/theta) T^-1 V, P2 = -(1/theta) T^-1 U."""
import numpy as np, sys
sys.path.insert(0, '/root/repo'); sys.path.insert(0, '/root/repo/scratch')
from oracle import fvm_oracle as O
from fvm_b200 import workloads as W
import proto_babe as PB
import proto_pipe as PP
L, C, N, M = 32, 32, 31, 16

def setup4(alo, adi, aup, theta, beta0=0.0):
    cf = PB.setup(alo, adi, aup, theta)
    alo = alo.reshape(L, C); adi = adi.reshape(L, C); aup = aup.reshape(L, C)
    l = np.zeros((L, N)); kap = np.zeros((L, N))
    sprev = np.zeros(L); invp = np.zeros(L)
    for c in range(M):
        inv = 1.0 / (adi[:, c] - alo[:, c] * sprev)
        l[:, c] = alo[:, c] * invp; kap[:, c] = inv / theta
        sprev = aup[:, c] * inv; invp = inv
    sprev = np.zeros(L); invp = np.zeros(L)
    for c in range(N-1, M-1, -1):
        inv = 1.0 / (adi[:, c] - aup[:, c] * sprev)
        l[:, c] = aup[:, c] * invp; kap[:, c] = inv / theta
        sprev = alo[:, c] * inv; invp = inv
    cf['l'], cf['kap'] = l, kap
    cf['omL'] = (1 - theta) * alo[:, 0]; cf['omR'] = (1 - theta) * aup[:, N-1]
    # P1, P2
    def solve_y(R):          # (1/theta) T^-1 R
        s, d = cf['s'], cf['delta']
        h = R.copy()
        for c in range(1, M): h[:, c] -= l[:, c] * h[:, c-1]
        for c in range(N-2, M-1, -1): h[:, c] -= l[:, c] * h[:, c+1]
        y = h * kap
        y[:, M] = (y[:, M] - s[:, M] * y[:, M-1]) * d
        for c in range(M-1, -1, -1): y[:, c] -= s[:, c] * y[:, c+1]
        for c in range(M+1, N): y[:, c] -= s[:, c] * y[:, c-1]
        return y
    cf['P1'] = -solve_y(cf['v']); cf['P2'] = -solve_y(cf['u'])
    return cf

def sweeps4(cf, g, outL, outR, bfirst=0.0):
    """(1/theta) T^-1 (g + coupling) - g, in place semantics of the kernel."""
    l, kap, s, d = cf['l'], cf['kap'], cf['s'], cf['delta']
    w = g.copy(); w_m_old = w[:, M].copy()
    hL = hR = hkL = hkR = None
    for i in range(M):
        c = i
        gc = w[:, c].copy(); gn = w_m_old if c == M-1 else w[:, c+1]
        hL = (gc - cf['omL'] * outL + bfirst) if i == 0 else gc - l[:, c] * hL
        if c == M-1: hkL = hL * kap[:, c]
        w[:, c] = hL * kap[:, c] - gc - s[:, c] * gn
        if i < N - M:
            c = N-1-i
            gc = w[:, c].copy(); gp = w[:, c-1]
            hR = (gc - cf['omR'] * outR) if i == 0 else gc - l[:, c] * hR
            if c == M: hkR = hR * kap[:, c]
            w[:, c] = hR * kap[:, c] - gc - s[:, c] * gp
    w[:, M] = (hkR - s[:, M] * hkL) * d - w_m_old
    for c in range(M-1, -1, -1): w[:, c] -= s[:, c] * w[:, c+1]
    for c in range(M+1, N): w[:, c] -= s[:, c] * w[:, c-1]
    return w

def run4(cf, w0, nsteps, theta, beta0=0.0):
    w = w0.reshape(L, C)
    nxt = lambda x: np.concatenate([x[1:], [0.0]]); prv = lambda x: np.concatenate([[0.0], x[:-1]])
    bf = np.zeros(L); bf[0] = theta * beta0
    g = sweeps4(cf, w[:, :N], prv(w[:, C-1]), w[:, C-1], bf)
    rs = cf['sblo'] * w[:, N-1] + cf['sbdi'] * w[:, C-1] + cf['sbup'] * nxt(w[:, 0])
    rr = rs - cf['slo'] * g[:, N-1] - cf['sup'] * nxt(g[:, 0])
    z = np.zeros(L)
    for n in range(1, nsteps):
        g0_old, gl_old = g[:, 0].copy(), g[:, N-1].copy()
        xs = PP.pcr(cf, rr)
        G = sweeps4(cf, g, z, z, bf)
        xsl = prv(xs)
        x30 = gl_old - cf['v'][:, N-1] * xsl - cf['u'][:, N-1] * xs
        x0 = g0_old - cf['v'][:, 0] * xsl - cf['u'][:, 0] * xs
        rs = cf['sblo'] * x30 + cf['sbdi'] * xs + cf['sbup'] * nxt(x0)
        g = G + cf['P1'] * xsl[:, None] + cf['P2'] * xs[:, None]
        rr = rs - cf['slo'] * g[:, N-1] - cf['sup'] * nxt(g[:, 0])
    xs = PP.pcr(cf, rr); xsl = prv(xs)
    x = np.zeros((L, C)); x[:, :N] = g - cf['v'] * xsl[:, None] - cf['u'] * xs[:, None]; x[:, C-1] = xs
    return x.reshape(-1)

if __name__ == '__main__':
    wl = W.batched_sweep(65536, 1024, count=3)
    theta = 0.5
    for p in range(3):
        asm = O.assemble(wl['faces'], wl['a'][p], wl['d'][p], 'exponential', **wl['bc'])
        A, B = O.theta_operators(asm, wl['k'][p], theta)
        cf = setup4(A['lo'], A['di'], A['up'], theta)
        for ns in (1, 2, 300):
            ref = O.theta_run_scipy(asm, wl['k'][p], theta, wl['w0'][p], ns, every=ns)[0]
            got = run4(cf, wl['w0'][p], ns, theta)
            print('cfg3 p', p, 'nsteps', ns, np.max(np.abs(got - ref)) / np.max(np.abs(ref)))
    # nonzero left flux (beta at the first cell), ragged J with padding
    J = 1000; f = W.graded_faces(J, 10.0); c = 0.5*(f[1:]+f[:-1]); w0 = W.pulse(c) + 0.2
    bc = dict(left_flux=0.3, right_value=0.0)
    asm = O.assemble(f, 0.8, 0.01, 'upwind', **bc)
    A, B = O.theta_operators(asm, 2e-3, theta)
    pad = lambda x, fill: np.concatenate([x, np.full(1024 - J, fill)])
    cf = setup4(pad(A['lo'], 0.), pad(A['di'], 1.), pad(A['up'], 0.), theta)
    for ns in (1, 3, 80):
        ref = O.theta_run_scipy(asm, 2e-3, theta, w0, ns, every=ns)[0]
        got = run4(cf, pad(w0, 0.0), ns, theta, beta0=asm['beta'][0])[:J]
        print('flux/ragged nsteps', ns, np.max(np.abs(got - ref)) / np.max(np.abs(ref)), 'alpha0', asm['alpha'][0])
