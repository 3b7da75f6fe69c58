"""Lane-level NumPy emulation of mibminet_b200/csrc/mibmi_fused.cu (test infrastructure).

It executes the same index arithmetic as the CUDA kernel -- fragment ownership of mma.sync
m16n8k32 / m16n8k16, ldmatrix row addressing, the PRMT byte transpose, the Toeplitz copies, the
magic-number division, the saturating pack -- on the packed network the library's host code
produces (mibmi_debug_pack_fused), so that packing / indexing bugs are caught on the CPU.  What it
cannot check is the hardware itself; tests/test_parity_gpu.py does that.
"""
import numpy as np

LANES = np.arange(32)
G, Q = LANES >> 2, LANES & 3
TOEP_COPY, TOEP_CHAN = 192, 768

FUSED_DTYPE = np.dtype({
    "names": ["w1frag", "off1", "m1", "sh1", "w2sh", "thr2", "sig2", "soff2", "m2", "sh2", "w3sh", "m3", "sh3",
              "w4frag", "thr4", "sig4", "soff4", "m4", "sh4", "w5p", "b5", "m5", "sh5"],
    "formats": [("<u4", 256), ("<i4", 16), ("<i4", 16), ("<i4", 16), ("u1", 16 * TOEP_CHAN),
                ("<i4", 16), ("<i4", 16), ("<i4", 16), ("<i4", 16), ("<i4", 16), ("<u4", 320), "<i4", "<i4",
                ("<u4", 64), ("<i4", 16), ("<i4", 16), ("<i4", 16), ("<i4", 16), ("<i4", 16),
                ("i1", 2048), ("<i4", 4), "<i4", "<i4"],
    "offsets": [0, 1024, 1088, 1152, 1216, 13504, 13568, 13632, 13696, 13760, 13824, 15104, 15108,
                15112, 15368, 15432, 15496, 15560, 15624, 15696, 17744, 17760, 17764],
    "itemsize": 17776})


def parse(blob: bytes):
    assert len(blob) == FUSED_DTYPE.itemsize, len(blob)
    return np.frombuffer(blob, dtype=FUSED_DTYPE)[0]


def _bytes(words):  # uint32 [...] -> int8 [..., 4] little endian
    w = np.asarray(words, dtype="<u4")
    return np.ascontiguousarray(w).view(np.int8).reshape(w.shape + (4,))


def _word(b):  # int8 [..., 4] -> uint32 [...]
    return np.ascontiguousarray(np.asarray(b, dtype=np.int8)).view("<u4").reshape(np.asarray(b).shape[:-1])


def mma_k32(acc, a, b0, b1):
    """acc int64 [32 lanes][4]; a uint32 [32][4]; b0, b1 uint32 [32] (PTX ISA: m16n8k32 .s8 fragments)."""
    A = np.zeros((16, 32), np.int64)
    Bm = np.zeros((32, 8), np.int64)
    ab = _bytes(a)
    for ln in range(32):
        g, q = ln >> 2, ln & 3
        A[g, 4 * q:4 * q + 4] = ab[ln, 0]
        A[g + 8, 4 * q:4 * q + 4] = ab[ln, 1]
        A[g, 16 + 4 * q:16 + 4 * q + 4] = ab[ln, 2]
        A[g + 8, 16 + 4 * q:16 + 4 * q + 4] = ab[ln, 3]
        Bm[4 * q:4 * q + 4, g] = _bytes(b0[ln])
        Bm[16 + 4 * q:16 + 4 * q + 4, g] = _bytes(b1[ln])
    D = A @ Bm
    out = acc.copy()
    out[:, 0] += D[G, 2 * Q]
    out[:, 1] += D[G, 2 * Q + 1]
    out[:, 2] += D[G + 8, 2 * Q]
    out[:, 3] += D[G + 8, 2 * Q + 1]
    return out


def mma_k16(a0, a1, b0):
    A = np.zeros((16, 16), np.int64)
    Bm = np.zeros((16, 8), np.int64)
    for ln in range(32):
        g, q = ln >> 2, ln & 3
        A[g, 4 * q:4 * q + 4] = _bytes(a0[ln])
        A[g + 8, 4 * q:4 * q + 4] = _bytes(a1[ln])
        Bm[4 * q:4 * q + 4, g] = _bytes(b0[ln])
    D = A @ Bm
    return np.stack([D[G, 2 * Q], D[G, 2 * Q + 1], D[G + 8, 2 * Q], D[G + 8, 2 * Q + 1]], axis=1)


def tdiv_magic(x, m, s):
    x = np.asarray(x, np.int64)
    assert np.all(np.abs(x) < 2 ** 31)
    return ((x * np.asarray(m, np.int64)) >> 32 >> np.asarray(s, np.int64)) + (x < 0)


def clip8(x):
    return np.clip(x, -128, 127)


def byte_perm(a, b, sel):
    src = np.concatenate([_bytes(a), _bytes(b)], axis=-1)  # [..., 8]
    idx = [(sel >> (4 * i)) & 7 for i in range(4)]
    return _word(np.stack([src[..., i] for i in idx], axis=-1))


def forward_trial(net, C, T, x):
    """x int8 [C, T] -> (logits int8 [4], taps) following the kernel's data flow for one trial."""
    KSTEPS = (C + 31) // 32
    T8, T64 = T // 8, T // 64
    UNITS, MROWS = (T + 31) // 32, (T8 * 8 + 31) // 32
    MT = (MROWS + 15) // 16
    y1need = max(32 + 32 * UNITS, 32 * (MROWS - 1) + 96)
    Y1S = ((y1need - 32 + 127) // 128) * 128 + 32
    Y2S = ((T8 + 24 + 15) // 16) * 16
    T64P = (T64 + 3) // 4 * 4
    XROWS = 32 * KSTEPS
    xs = np.zeros(XROWS * T + 128, np.int8)
    xs[:C * T] = x.ravel()
    xw = xs[:(XROWS * T // 4) * 4].view("<u4")
    y1 = np.zeros(16 * Y1S + 128, np.int8)
    y2 = np.zeros(16 * Y2S, np.int8)
    y3 = np.zeros(((T8 + 7) // 8 * 8) * 16, np.int8)
    y4 = np.zeros(16 * T64P, np.int8)

    w1 = net["w1frag"]
    afrag = [np.stack([w1[((s * 2 + 0) * 16 + G) * 4 + Q], w1[((s * 2 + 0) * 16 + G + 8) * 4 + Q],
                       w1[((s * 2 + 1) * 16 + G) * 4 + Q], w1[((s * 2 + 1) * 16 + G + 8) * 4 + Q]], axis=1)
             for s in range(KSTEPS)]
    # ---- layer 1
    for u in range(UNITS):
        acc = [np.stack([net["off1"][G], net["off1"][G], net["off1"][G + 8], net["off1"][G + 8]], axis=1).astype(np.int64)
               for _ in range(4)]
        for s in range(KSTEPS):
            bf = [[None] * 4, [None] * 4]
            for h in range(2):
                base = (32 * s + 16 * h + Q) * (T // 4) + 8 * u + G
                r0, r1, r2, r3 = (xw[base + 4 * i * (T // 4)] for i in range(4))
                t0, t1 = byte_perm(r0, r1, 0x5140), byte_perm(r2, r3, 0x5140)
                t2, t3 = byte_perm(r0, r1, 0x7362), byte_perm(r2, r3, 0x7362)
                bf[h][0], bf[h][1] = byte_perm(t0, t1, 0x5410), byte_perm(t0, t1, 0x7632)
                bf[h][2], bf[h][3] = byte_perm(t2, t3, 0x5410), byte_perm(t2, t3, 0x7632)
            for j in range(4):
                acc[j] = mma_k32(acc[j], afrag[s], bf[0][j], bf[1][j])
        for half, (mm, ss) in enumerate(((net["m1"][G], net["sh1"][G]), (net["m1"][G + 8], net["sh1"][G + 8]))):
            for c in range(2):
                vals = np.stack([clip8(tdiv_magic(acc[j][:, 2 * half + c], mm, ss)) for j in range(4)], axis=1)
                for ln in range(32):
                    g, q = ln >> 2, ln & 3
                    row = g + 8 * half
                    t0 = 32 * u + 8 * q + 4 * c
                    for i in range(4):
                        y1[row * Y1S + 32 + t0 + i] = vals[ln, i] if t0 + i < T else 0
    tap1 = np.stack([y1[k * Y1S + 32:k * Y1S + 32 + T] for k in range(16)])
    # ---- layers 2 + 3
    toep_base = []
    for v in range(4):
        o = 8 * (G >> 1) + 2 * v + (G & 1)
        j0 = 31 - o
        toep_base.append((j0 & 3) * TOEP_COPY + (j0 & ~3) + 4 * Q)
    for k in range(16):
        tp = net["w2sh"][k * TOEP_CHAN:(k + 1) * TOEP_CHAN]
        tpw = lambda off: _word(np.stack([tp[off + i] for i in range(4)], axis=-1).astype(np.int8))
        bt = [[[tpw(toep_base[v] + 32 * s + 16 * h) for h in range(2)] for s in range(3)] for v in range(4)]
        for mt in range(MT):
            row = np.minimum(16 * mt + (LANES & 7) + 8 * ((LANES >> 3) & 1), MROWS - 1)
            abase = k * Y1S + 32 * row + 16 * (LANES >> 4)
            acc = [np.zeros((32, 4), np.int64) for _ in range(4)]
            for s in range(3):
                # ldmatrix.x4: lane l supplies the row address of row (l & 7) of matrix (l >> 3);
                # matrix i lands in register i, thread (g, q) gets bytes 4q..4q+3 of row g
                a = np.zeros((32, 4), "<u4")
                for mi in range(4):
                    for ln in range(32):
                        g, q = ln >> 2, ln & 3
                        addr = abase[mi * 8 + g] + 32 * s + 4 * q
                        a[ln, mi] = _word(y1[addr:addr + 4])
                for v in range(4):
                    acc[v] = mma_k32(acc[v], a, bt[v][s][0], bt[v][s][1])
            for hf in range(2):
                ssum = sum(np.maximum(acc[v][:, 2 * hf], net["thr2"][k]) + np.maximum(acc[v][:, 2 * hf + 1], net["thr2"][k])
                           for v in range(4))
                p = 64 * mt + 32 * hf + LANES
                o = clip8(tdiv_magic(ssum * net["sig2"][k] + net["soff2"][k], net["m2"][k], net["sh2"][k]))
                for ln in range(32):
                    if p[ln] < T8:
                        y2[k * Y2S + 8 + p[ln]] = o[ln]
        y2w = y2[k * Y2S:(k + 1) * Y2S].view("<u4")
        for t in range(T8):
            r = (t + 1) & 3
            w3 = net["w3sh"][(k * 4 + r) * 5:(k * 4 + r) * 5 + 5]
            a = 0
            for i in range(5):
                a += int(np.dot(_bytes(y2w[((t + 1) >> 2) + i]).astype(np.int64), _bytes(w3[i]).astype(np.int64)))
            y3[t * 16 + k] = clip8(tdiv_magic(a, net["m3"], net["sh3"]))
    tap2 = np.stack([y2[k * Y2S + 8:k * Y2S + 8 + T8] for k in range(16)])
    tap3 = y3[:T8 * 16].reshape(T8, 16).T.copy()
    # ---- layer 4
    y3w = y3.view("<u4")
    for j in range(T64):
        acc = mma_k16(net["w4frag"][G * 4 + Q], net["w4frag"][(G + 8) * 4 + Q], y3w[(8 * j + G) * 4 + Q])
        tl, th = net["thr4"][G], net["thr4"][G + 8]
        s_lo = np.maximum(acc[:, 0], tl) + np.maximum(acc[:, 1], tl)
        s_hi = np.maximum(acc[:, 2], th) + np.maximum(acc[:, 3], th)
        for xor in (1, 2):
            s_lo = s_lo + s_lo[LANES ^ xor]
            s_hi = s_hi + s_hi[LANES ^ xor]
        for ln in range(0, 32, 4):
            g = ln >> 2
            y4[g * T64P + j] = clip8(tdiv_magic(s_lo[ln] * net["sig4"][g] + net["soff4"][g], net["m4"][g], net["sh4"][g]))
            y4[(g + 8) * T64P + j] = clip8(tdiv_magic(s_hi[ln] * net["sig4"][g + 8] + net["soff4"][g + 8],
                                                      net["m4"][g + 8], net["sh4"][g + 8]))
    tap4 = y4.reshape(16, T64P)[:, :T64].copy()
    # ---- layer 5
    w5 = net["w5p"].reshape(4, 16 * 32)[:, :16 * T64P] if T64P == 32 else net["w5p"][:4 * 16 * T64P].reshape(4, 16 * T64P)
    z = w5.astype(np.int64) @ y4.astype(np.int64) + net["b5"]
    out = clip8(tdiv_magic(z, net["m5"], net["sh5"])).astype(np.int8)
    return out, {"y1": tap1, "y2": tap2, "y3": tap3, "y4": tap4}
