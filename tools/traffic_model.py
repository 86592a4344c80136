"""Back-of-envelope per-layer bounds for the ResNet-50 fwd+dgrad convs at batch N (bf16):
tensor time, HBM time, and L2<->SM fabric time for a given tiling policy."""
import sys
N = int(sys.argv[1]) if len(sys.argv) > 1 else 256
PTF, PBW, PL2 = 1680e12, 6450e9, 15e12   # flop/s, HBM B/s, L2<->SM B/s (6300 B/clk chip-wide)

layers = []  # (name, M per img, K(cin), N(cout), taps, count, n_aux_fwd, n_aux_bwd)
def blk(l, hin, cin, w, stride, nb):
    hout = hin // stride
    for b in range(nb):
        ci = cin if b == 0 else 4 * w
        hi = hin if b == 0 else hout
        layers.append((f"l{l}.{b}.c1", hi * hi, ci, w, 1, 0))
        layers.append((f"l{l}.{b}.c2", hout * hout, w, w, 9, 0))
        layers.append((f"l{l}.{b}.c3", hout * hout, w, 4 * w, 1, 1))
        if b == 0:
            layers.append((f"l{l}.{b}.ds", hout * hout, ci, 4 * w, 1, 0))
blk(1, 56, 64, 64, 1, 3); blk(2, 56, 256, 128, 2, 4); blk(3, 28, 512, 256, 2, 6); blk(4, 14, 1024, 512, 2, 3)

def bound(M, K, Nn, taps, naux, BM, BN, a_reuse, b_share):
    """times (s) for one GEMM: M x Nn outputs, K*taps reduction."""
    fl = 2.0 * M * Nn * K * taps
    hbm = 2.0 * (M * K + M * Nn * (1 + naux) + taps * K * Nn)
    a = 2.0 * M * K * taps / a_reuse * (Nn / min(BN, Nn))
    b = 2.0 * (M / BM) * K * taps * Nn * b_share
    epi = 2.0 * M * Nn * (1 + naux)
    return fl / PTF, hbm / PBW, (a + b + epi) / PL2

tot = {"cur": 0.0, "new": 0.0, "hbm": 0.0, "tc": 0.0}
for name, m, k, n, taps, aux in layers:
    M = m * N
    for d, (kk, nn, ax) in (("fwd", (k, n, aux)), ("bwd", (n, k, 2 if name.endswith("c1") else (0 if name.endswith("ds") else 1)))):
        bn_cur = 64 if nn == 64 else 128
        t_tc, t_hbm, t_l2 = bound(M, kk, nn, taps, ax, 128, bn_cur, 1.0, 1.0)
        cur = max(t_tc, t_hbm, t_l2)
        # new policy: 2-CTA 256 x min(N,256) tiles (B split across the pair), 3x3 A rows shared across dh (2.4x)
        t_tc2, t_hbm2, t_l22 = bound(M, kk, nn, taps, ax, 256, 256, 2.4 if taps == 9 else 1.0, 1.0)
        new = max(t_tc2, t_hbm2, t_l22)
        tot["cur"] += cur; tot["new"] += new; tot["hbm"] += t_hbm; tot["tc"] += t_tc
        print(f"{d} {name:10s} tc {t_tc*1e6:6.1f} hbm {t_hbm*1e6:6.1f} l2cur {t_l2*1e6:6.1f} l2new {t_l22*1e6:6.1f} -> cur {cur*1e6:6.1f} new {new*1e6:6.1f}")
print({k: round(v * 1e6) for k, v in tot.items()})
