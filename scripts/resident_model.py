"""Lane-level numpy model of csrc/cuda/stencil_resident.cu (no GPU needed).

Every register of the kernel is an array [cta, warp, lane] of uint32 and every statement of
resident_kernel / res_iteration / hsum is repeated on those arrays in the same order, including the
things that only matter at the edges: SHFL returning the lane's own value at lane 0 / 31, rows
above warp 0 and below the last warp contributing zeros, dead cells, the scale-1 / scale-16
alternation, the in-place row order, the ring / halo classification and the exchange images.
What it cannot show is the CUDA side (barriers, memory ordering, co-residency).

    python scripts/resident_model.py          # compares a set of shapes with the CPU oracle
"""
import sys
from pathlib import Path

import numpy as np

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))

WARPS, K, NW, SB, WINB = 16, 4, 2, 8, 256
U = np.uint32


def halo_bytes(ch):
    return (K * ch + SB - 1) // SB * SB


def plan(bw, H, ch, sms=148):
    HB = halo_bytes(ch)
    VC = WINB - 2 * HB
    if bw <= 0 or H < K or bw < HB or bw % SB:
        return None
    strips = -(-bw // VC)
    if bw - (strips - 1) * VC < HB:
        return None
    for rw in (4, 6, 8, 10, 12, 14, 16):
        VR = WARPS * rw - 2 * K
        ctarows = -(-H // VR)
        if H - (ctarows - 1) * VR < K:
            continue
        if strips * ctarows > sms:
            continue
        return dict(rw=rw, strips=strips, ctarows=ctarows, vr=VR, vc=VC)
    return None


def funnel_l(lo, hi, sh):   # __funnelshift_l(lo, hi, 16): upper 32 bits of (hi:lo) << 16
    return ((hi << U(sh)) | (lo >> U(32 - sh))).astype(U)


def byte_perm(a, b, sel):
    a = a.astype(np.uint64)
    b = np.broadcast_to(np.asarray(b, dtype=np.uint64), a.shape)
    both = a | (b << np.uint64(32))
    sel = np.broadcast_to(np.asarray(sel, dtype=np.uint32), a.shape)
    out = np.zeros(a.shape, dtype=np.uint64)
    for i in range(4):
        n = ((sel >> U(4 * i)) & U(7)).astype(np.uint64)
        out |= ((both >> (n * np.uint64(8))) & np.uint64(0xFF)) << np.uint64(8 * i)
    return out.astype(U)


def shfl_down(x):   # delta 1; the last lane keeps its own value
    y = x.copy()
    y[..., :-1] = x[..., 1:]
    return y


def shfl_up(x):
    y = x.copy()
    y[..., 1:] = x[..., :-1]
    return y


def hsum(E, O, ch):
    """E, O: lists of NW arrays [..., lane].  Mirrors hsum<NW,CH> in stencil_stream.cuh."""
    tE, tO, hE, hO = [None] * NW, [None] * NW, [None] * NW, [None] * NW
    if ch == 1:
        En = shfl_down(E[0])
        for i in range(NW):
            tE[i] = (E[i] + O[i]).astype(U)
            tO[i] = (funnel_l(E[i], E[i + 1] if i + 1 < NW else En, 16) + O[i]).astype(U)
        tOp = shfl_up(tO[NW - 1])
        for i in range(NW):
            hE[i] = (funnel_l(tO[i - 1] if i else tOp, tO[i], 16) + tE[i]).astype(U)
            hO[i] = (tE[i] + tO[i]).astype(U)
    else:
        On, En = shfl_down(O[0]), shfl_down(E[0])
        for i in range(NW):
            tE[i] = (funnel_l(O[i], O[i + 1] if i + 1 < NW else On, 16) + E[i]).astype(U)
            tO[i] = ((E[i + 1] if i + 1 < NW else En) + O[i]).astype(U)
        tOp, tEp = shfl_up(tO[NW - 1]), shfl_up(tE[NW - 1])
        for i in range(NW):
            hE[i] = ((tO[i - 1] if i else tOp) + tE[i]).astype(U)
            hO[i] = (funnel_l(tE[i - 1] if i else tEp, tE[i], 16) + tO[i]).astype(U)
    return hE, hO


def run(img, reps, boundary, sms=148):
    """img: uint8 [H, bw] (bytes).  Returns the blurred bytes the kernel would write."""
    H, bw = img.shape
    ch = run.ch
    torus = boundary == "torus"
    pl = plan(bw, H, ch, sms)
    assert pl, "image does not fit"
    RW, strips, ctarows, VR, VC = pl["rw"], pl["strips"], pl["ctarows"], pl["vr"], pl["vc"]
    HB = halo_bytes(ch)
    ncta = strips * ctarows
    cta = np.arange(ncta)
    bx, by = (cta % strips)[:, None, None], (cta // strips)[:, None, None]
    warp = np.arange(WARPS)[None, :, None]
    lane = np.arange(32)[None, None, :]
    vr0 = by * VR
    vr1 = np.minimum(vr0 + VR, H)
    vc0 = bx * VC
    vc1 = np.minimum(vc0 + VC, bw)
    gc = vc0 - HB + lane * SB + 0 * warp
    wr0 = vr0 - K + warp * RW + 0 * lane
    col_valid = (gc >= vc0) & (gc < vc1)
    col_near = (gc >= vc0 - HB) & (gc < vc1 + HB)
    col_ring = col_valid & ((gc < vc0 + HB) | (gc >= vc1 - HB))
    lane_ok = np.ones_like(gc, bool) if torus else ((gc >= 0) & (gc < bw))

    def load8(image, gr, gcol, want):
        """-> two uint32 word arrays; zero where `want` is false or outside (zero mode)."""
        gr, gcol = gr.copy(), gcol.copy()
        if torus:
            gr = np.where(gr < 0, gr + H, np.where(gr >= H, gr - H, gr))
            gcol = np.where(gcol < 0, gcol + bw, np.where(gcol >= bw, gcol - bw, gcol))
        ok = want & (gr >= 0) & (gr < H) & (gcol >= 0) & (gcol < bw)
        grc, gcc = np.where(ok, gr, 0), np.where(ok, gcol, 0)
        words = []
        for w in range(2):
            v = np.zeros(gr.shape, dtype=U)
            for b in range(4):
                v |= image[grc, gcc + 4 * w + b].astype(U) << U(8 * b)
            words.append(np.where(ok, v, U(0)).astype(U))
        return words

    def unpack(words):
        return ([(words[i] & U(0x00FF00FF)).astype(U) for i in range(2)],
                [byte_perm(words[i], 0, 0x4341) for i in range(2)])

    def pack(E, O):
        return [byte_perm(E[i], O[i], 0x6240) for i in range(2)]

    def store8(image, gr, gcol, words, want):
        idx = np.nonzero(want)
        for w in range(2):
            for b in range(4):
                image[gr[idx], gcol[idx] + 4 * w + b] = ((words[w][idx] >> U(8 * b)) & U(0xFF)).astype(np.uint8)

    E = [None] * RW
    O = [None] * RW
    for r in range(RW):
        gr = wr0 + r
        row_near = (gr >= vr0 - K) & (gr < vr1 + K)
        E[r], O[r] = unpack(load8(img, gr, gc, row_near & col_near))

    def iteration(odd):
        hc = hsum(E[0], O[0], ch)
        hL = hsum(E[RW - 1], O[RW - 1], ch)
        hcE, hcO = list(hc[0]), list(hc[1])
        psE, psO, dnE, dnO = [], [], [], []
        for i in range(NW):
            upE, upO = np.zeros_like(hcE[i]), np.zeros_like(hcO[i])
            upE[:, 1:, :], upO[:, 1:, :] = hL[0][i][:, :-1, :], hL[1][i][:, :-1, :]   # warp-1's last row
            psE.append((upE + hcE[i]).astype(U))
            psO.append((upO + hcO[i]).astype(U))
            dE, dO = np.zeros_like(hcE[i]), np.zeros_like(hcO[i])
            dE[:, :-1, :], dO[:, :-1, :] = hc[0][i][:, 1:, :], hc[1][i][:, 1:, :]     # warp+1's first row
            dnE.append(dE)
            dnO.append(dO)
        for r in range(RW):
            if r + 1 < RW - 1:
                hnE, hnO = hsum(E[r + 1], O[r + 1], ch)
            elif r + 1 == RW - 1:
                hnE, hnO = hL
            else:
                hnE, hnO = dnE, dnO
            gr = wr0 + r
            ok = lane_ok & (np.ones_like(gr, bool) if torus else ((gr >= 0) & (gr < H)))
            m = np.where(ok, U(0x0FF00FF0), U(0)).astype(U)
            sel = np.where(ok, U(0x4341), U(0x4444)).astype(U)
            newE, newO = [], []
            for i in range(NW):
                sE, sO = (hcE[i] + hnE[i]).astype(U), (hcO[i] + hnO[i]).astype(U)
                vE, vO = (psE[i] + sE).astype(U), (psO[i] + sO).astype(U)
                psE[i], psO[i], hcE[i], hcO[i] = sE, sO, hnE[i], hnO[i]
                newE.append(byte_perm(vE, 0, sel) if odd else (vE & m).astype(U))
                newO.append(byte_perm(vO, 0, sel) if odd else (vO & m).astype(U))
            E[r], O[r] = newE, newO

    xbuf = [np.zeros((H, bw), np.uint8), np.zeros((H, bw), np.uint8)]
    epoch, done = 0, 0
    while True:
        k = min(K, reps - done)
        it = 0
        while it < k:
            iteration(False)
            if it + 1 < k:
                iteration(True)
            it += 2
        if k & 1:
            for r in range(RW):
                E[r] = [(e >> U(4)).astype(U) for e in E[r]]
                O[r] = [(o >> U(4)).astype(U) for o in O[r]]
        done += k
        if done >= reps:
            break
        epoch += 1
        xb = xbuf[epoch & 1]
        xb[...] = 0xCD     # poison: every byte a halo needs must have been written in this epoch
        for r in range(RW):
            gr = wr0 + r
            row_valid = (gr >= vr0) & (gr < vr1)
            row_ring = (gr < vr0 + K) | (gr >= vr1 - K)
            store8(xb, gr, gc, pack(E[r], O[r]), row_valid & col_valid & (row_ring | col_ring))
        for r in range(RW):
            gr = wr0 + r
            row_valid = (gr >= vr0) & (gr < vr1)
            row_near = (gr >= vr0 - K) & (gr < vr1 + K)
            want = ~(row_valid & col_valid) & row_near & col_near
            nE, nO = unpack(load8(xb, gr, gc, want))
            for i in range(NW):
                E[r][i] = np.where(want, nE[i], E[r][i]).astype(U)
                O[r][i] = np.where(want, nO[i], O[r][i]).astype(U)
    out = np.zeros((H, bw), np.uint8)
    for r in range(RW):
        gr = wr0 + r
        store8(out, gr, gc, pack(E[r], O[r]), (gr >= vr0) & (gr < vr1) & col_valid)
    return out, pl


def main():
    from b200conv import reference_oracle
    rng = np.random.default_rng(0)
    cases = [  # (H, W, ch, reps, boundary, sms)
        (40, 64, 1, 9, "zero", 148), (40, 64, 1, 9, "torus", 148), (64, 256, 1, 8, "zero", 148),
        (130, 488, 1, 13, "torus", 148), (130, 488, 1, 13, "zero", 148), (200, 96, 3, 10, "zero", 148),
        (200, 96, 3, 11, "torus", 148), (77, 160, 3, 5, "zero", 148), (300, 520, 1, 6, "torus", 12),
        (61, 8, 1, 7, "torus", 148), (4, 16, 1, 3, "zero", 148), (120, 1920, 1, 12, "zero", 148),
    ]
    bad = 0
    for (H, W, ch, reps, boundary, sms) in cases:
        img = rng.integers(0, 256, (H, W) if ch == 1 else (H, W, ch), dtype=np.uint8)
        run.ch = ch
        if plan(W * ch, H, ch, sms) is None:
            print(f"{H}x{W}x{ch}: does not fit (plan rejects)")
            continue
        got, pl = run(img.reshape(H, W * ch), reps, boundary, sms)
        want = reference_oracle(img, reps, boundary).reshape(H, W * ch)
        ok = bool((got == want).all())
        bad += not ok
        print(f"{H}x{W}x{ch} reps={reps} {boundary} plan={pl} {'OK' if ok else 'MISMATCH %d bytes' % int((got != want).sum())}")
    print("bad =", bad)
    return bad


if __name__ == "__main__":
    sys.exit(1 if main() else 0)
