"""numpy prototype of the delayed-correction pipeline on top of proto_babe's twisted local solve:
state between steps = (g: block-local solutions, rr: reduced rhs); sweeps of step n+1 run on g_n
while PCR_n produces xs_n; then g_{n+1} = T^-1 B_ii g_n + P1*xsl_n + P2*xs_n."""
import numpy as np, sys
sys.path.insert(0, '/root/repo'); sys.path.insert(0, '/root/repo/scratch')
from oracle import fvm_oracle as O
from fvm_b200 import workloads as W
import proto_babe as PB
L, C, N, M = 32, 32, 31, 16

def sweeps(cf, g_int, wl, wsep):
    """T^-1 of r~ built from interior values g_int (L,N) with given outside neighbours (L,) each."""
    e, s, b, cth = cf['e'], cf['s'], cf['b'], cf['cth']
    old = g_int
    g = np.zeros((L, N))
    for c in range(M):
        wm = wl if c == 0 else old[:, c-1]
        t = s[:, c] * old[:, c+1] + e[:, c] * wm
        r = b[:, c] * old[:, c] - cth * t
        g[:, c] = r if c == 0 else r - e[:, c] * g[:, c-1]
    for c in range(N-1, M-1, -1):
        wp = wsep if c == N-1 else old[:, c+1]
        t = s[:, c] * old[:, c-1] + e[:, c] * wp
        r = b[:, c] * old[:, c] - cth * t
        g[:, c] = r if c == N-1 else r - e[:, c] * g[:, c+1]
    g[:, M] = (g[:, M] - s[:, M] * g[:, M-1]) * cf['delta']
    for c in range(M-1, -1, -1): g[:, c] -= s[:, c] * g[:, c+1]
    for c in range(M+1, N): g[:, c] -= s[:, c] * g[:, c-1]
    return g

def pcr(cf, rr):
    st = 1
    for a_, g_ in zip(cf['al'], cf['ga']):
        rm = np.zeros(L); rm[st:] = rr[:-st]; rp = np.zeros(L); rp[:-st] = rr[st:]
        rr = rr + a_ * rm + g_ * rp; st *= 2
    return rr * cf['invD']

def setup_P(cf):
    e, s, b, cth, V, U = cf['e'], cf['s'], cf['b'], cf['cth'], cf['v'], cf['u']
    z = np.zeros(L)
    def Bii(X):   # r~ of interior vector X with zero outside
        R = np.zeros((L, N))
        for c in range(N):
            if c < M: en = X[:, c-1] if c > 0 else z; sn = X[:, c+1]
            else:     en = X[:, c+1] if c < N-1 else z; sn = X[:, c-1]
            R[:, c] = b[:, c] * X[:, c] - cth * (e[:, c] * en + s[:, c] * sn)
        return R
    def solve(R):
        g = R.copy()
        for c in range(1, M): g[:, c] -= e[:, c] * g[:, c-1]
        for c in range(N-2, M-1, -1): g[:, c] -= e[:, c] * g[:, c+1]
        g[:, M] = (g[:, M] - s[:, M] * g[:, M-1]) * cf['delta']
        for c in range(M-1, -1, -1): g[:, c] -= s[:, c] * g[:, c+1]
        for c in range(M+1, N): g[:, c] -= s[:, c] * g[:, c-1]
        return g
    R1 = -Bii(V); R1[:, 0] += -cth * e[:, 0]
    R2 = -Bii(U); R2[:, N-1] += -cth * e[:, N-1]
    cf['P1'], cf['P2'] = solve(R1), solve(R2)

def run(cf, w0, nsteps):
    w = w0.reshape(L, C)
    nxt = lambda x: np.concatenate([x[1:], [0.0]]); prv = lambda x: np.concatenate([[0.0], x[:-1]])
    # prologue: standard sweeps from the true state
    g = sweeps(cf, w[:, :N], prv(w[:, C-1]), w[:, C-1])
    rs = cf['sblo'] * w[:, N-1] + cf['sbdi'] * w[:, C-1] + cf['sbup'] * nxt(w[:, 0])
    rr = rs - cf['slo'] * g[:, N-1] - cf['sup'] * nxt(g[:, 0])
    for n in range(1, nsteps):
        g0_old, gl_old = g[:, 0].copy(), g[:, N-1].copy()
        xs = pcr(cf, rr)                                  # overlaps with ...
        G = sweeps(cf, g, np.zeros(L), np.zeros(L))        # ... the sweeps on the uncorrected g
        xsl = prv(xs)
        x30 = gl_old - cf['v'][:, N-1] * xsl - cf['u'][:, N-1] * xs
        x0 = g0_old - cf['v'][:, 0] * xsl - cf['u'][:, 0] * xs
        rs = cf['sblo'] * x30 + cf['sbdi'] * xs + cf['sbup'] * nxt(x0)
        g = G + cf['P1'] * xsl[:, None] + cf['P2'] * xs[:, None]
        rr = rs - cf['slo'] * g[:, N-1] - cf['sup'] * nxt(g[:, 0])
    xs = pcr(cf, rr); xsl = prv(xs)
    x = np.zeros((L, C)); x[:, :N] = g - cf['v'] * xsl[:, None] - cf['u'] * xs[:, None]; x[:, C-1] = xs
    return x.reshape(-1)

if __name__ == '__main__':
    wl = W.batched_sweep(65536, 1024, count=3)
    for theta in (0.5, 1.0, 0.7):
        for p in range(3):
            asm = O.assemble(wl['faces'], wl['a'][p], wl['d'][p], 'exponential', **wl['bc'])
            A, B = O.theta_operators(asm, wl['k'][p], theta)
            cf = PB.setup(A['lo'], A['di'], A['up'], theta); setup_P(cf)
            for ns in (1, 2, 300):
                ref = O.theta_run_scipy(asm, wl['k'][p], theta, wl['w0'][p], ns, every=ns)[0]
                got = run(cf, wl['w0'][p], ns)
                print('theta', theta, 'p', p, 'nsteps', ns, np.max(np.abs(got - ref)) / np.max(np.abs(ref)))
