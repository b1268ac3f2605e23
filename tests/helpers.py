"""Test helpers: bf16 rounding, synthetic inputs, and a numpy emulation of the
tap-GEMM plan (what the TMA boxes + step table + packed weights compute)."""
import numpy as np


def bf16_round(x):
    """Round float32 -> bfloat16 -> float32 (round to nearest even)."""
    x = np.ascontiguousarray(x, dtype=np.float32)
    u = x.view(np.uint32).astype(np.uint64)
    u = (u + 0x7FFF + ((u >> 16) & 1)) & 0xFFFF0000
    return u.astype(np.uint32).view(np.float32).reshape(x.shape)


def bf16_bits_to_f32(bits):
    return (np.asarray(bits, dtype=np.uint32) << 16).view(np.float32)


def exact16(x):
    """Test inputs exactly representable in BOTH operand formats: bf16-rounded, with magnitudes below fp16's smallest
    normal (2^-14) flushed to zero (8 significant bits always fit fp16's 11)."""
    r = bf16_round(x)
    return np.where(np.abs(r) < 2.0 ** -14, np.float32(0), r).astype(np.float32)


def fp16_round(x):
    """Round float -> float16 -> float32."""
    return np.asarray(x, dtype=np.float32).astype(np.float16).astype(np.float32)


def op16_round(x, dtype='fp16'):
    """Round through the library's 16-bit operand format (csrc/op16.h)."""
    return fp16_round(x) if dtype == 'fp16' else bf16_round(x)


def op16_bits_to_f32(bits, dtype='fp16'):
    """Packed 16-bit operand words (as the planner emits them) -> float32."""
    if dtype == 'fp16':
        return np.ascontiguousarray(bits, dtype=np.uint16).view(np.float16).astype(np.float32)
    return bf16_bits_to_f32(bits)


def pad_channels(x, cp):
    out = np.zeros(x.shape[:-1] + (cp,), dtype=x.dtype)
    out[..., :x.shape[-1]] = x
    return out


def emulate_plan(plan, srcs, n_valid):
    """Evaluate a Plan on the CPU exactly as the kernel would: for every tile
    and step, gather the 5-D TMA box (zero fill out of bounds) from the padded
    NHWC source, multiply by the packed weight slice, accumulate; add bias.
    srcs: list of float arrays [N, H, W, Cp].  Returns [n_valid, Ho, Wo, npad]."""
    i = plan.info
    steps, wts, bias = plan.steps, op16_bits_to_f32(plan.weights, plan.dtype).astype(np.float64), plan.bias
    out = np.zeros((n_valid, i.ho, i.wo, i.npad), dtype=np.float64)
    views = []
    for s in range(i.nsrc):
        dims = [int(v) for v in i.src_dims[s]]
        strides = [int(v) for v in i.src_strides[s]]
        flat = np.ascontiguousarray(srcs[s], dtype=np.float64).ravel()
        # element strides of the 5-D view (bf16 = 2 bytes)
        est = [1] + [st // 2 for st in strides]
        views.append((flat, dims, est))

    def gather(s, c0, x0, p0, y0, n0):
        flat, dims, est = views[s]
        box = [int(v) for v in i.src_box[s]]
        a = np.zeros((box[4], box[3], box[2], box[1], box[0]), dtype=np.float64)
        for bn in range(box[4]):
            for by in range(box[3]):
                for bx in range(box[1]):
                    n, y, x, p = n0 + bn, y0 + by, x0 + bx, p0
                    if not (0 <= n < dims[4] and 0 <= y < dims[3] and 0 <= x < dims[1] and 0 <= p < dims[2]):
                        continue
                    base = n * est[4] + y * est[3] + p * est[2] + x * est[1]
                    for bc in range(box[0]):
                        c = c0 + bc
                        if 0 <= c < dims[0]:
                            a[bn, by, 0, bx, bc] = flat[base + c]
        return a.reshape(-1, box[0])      # rows ordered (n, y, x) = kernel row order

    tiles_n = -(-n_valid // i.bn)
    for tn in range(tiles_n):
        for ty in range(i.tiles_y):
            for tx in range(i.tiles_x):
                n0, y0, x0 = tn * i.bn, ty * i.bh, tx * i.bw
                for ph in range(i.num_phases):
                    acc = np.zeros((i.rows, i.npad), dtype=np.float64)
                    for st in range(i.phase_begin[ph], i.phase_begin[ph + 1]):
                        src, dx, dy, p, c0 = [int(v) for v in steps[st]]
                        a = gather(src, c0, x0 + dx, p, y0 + dy, n0)
                        acc += a @ wts[st].T
                    acc += bias
                    py, px = (ph >> 1, ph & 1) if i.num_phases == 4 else (0, 0)
                    for r in range(i.rows):
                        x = r % i.bw
                        y = (r // i.bw) % i.bh
                        n = r // (i.bw * i.bh)
                        if n0 + n >= n_valid or y0 + y >= i.hg or x0 + x >= i.wg:
                            continue
                        out[n0 + n, (y0 + y) * i.os + py, (x0 + x) * i.os + px] = acc[r]
    return out


def synthetic_images(shape, seed=0):
    """uint8 ~ U{0..255} mapped as dmm_cifar10.py:1084-1085."""
    u8 = np.random.RandomState(seed).randint(0, 256, size=shape)
    return ((u8 - 127.5) / 256.0 * 2).astype(np.float32)


def emulate_patch_plan(plan, srcs, n_valid):
    """Evaluate the A-stationary ("patch") plan on the CPU exactly as patchgemm.cu would: per tile and
    group, one zero-filled patch [prow, W, kc]; tap j reads the patch advanced by j*W rows, accumulator
    m by 128 rows.  Stride-2 convs (patch_s2d) read the space-to-depth view of the input.  Folded plans (8-wide images, patch_fold == 2) lay two images side by side: the patch
    is [prow, 2 images, 8, kc] and a tile row is 16 pixels.  Virtual widths (patch_wv > W: 28 in 32, 14 in 16, 7 in 8):
    the tile rows are patch_wv wide, what hangs over the image's right / bottom edge is zero on load and dropped on
    store.  srcs: float arrays [N, H, W, Cp].  Returns [n_valid, H, W, npad]."""
    i = plan.info
    assert i.patch
    groups = plan.patch_groups
    wts = op16_bits_to_f32(plan.patch_weights, plan.dtype).astype(np.float64)
    npad = i.patch_npad or i.npad
    bias = plan.patch_bias if i.patch_npad else plan.bias
    H, W, kc, macc, rt = i.hg, i.wg, i.kc, i.patch_macc, i.patch_rows_tile
    fold = max(1, i.patch_fold)
    Wv = i.patch_wv or W                 # tile-row width per image
    Wt = Wv * fold
    # a sub-pixel phase plan (Plan.select_patch) writes output pixels (2y + py, 2x + px) of a 2H x 2W image; a row-phase
    # plan (phase 4 + py) carries both column phases side by side in N
    ph = i.patch_phase
    cout = i.cout
    out = np.zeros((n_valid, H, W, npad) if ph < 0 else (n_valid, 2 * H, 2 * W, cout), dtype=np.float64)
    tiles_y = -(-H // rt)
    for n0 in range(0, n_valid, fold):
        for ty in range(tiles_y):
            y0 = ty * rt
            acc = np.zeros((macc * 128, npad), dtype=np.float64)
            item = 0
            for mp, dx, py0, k, c0, prow in groups:
                x = np.asarray(srcs[i.patch_map_src[mp]], dtype=np.float64)
                if i.patch_s2d:      # space-to-depth view of the 2H x 2W input: [N, H, row phase, W, 2 Cp], one row phase per map
                    nn, hh, ww, cc = x.shape
                    x = x.reshape(nn, hh // 2, 2, ww // 2, 2 * cc)[:, :, i.patch_map_ph[mp]]
                patch = np.zeros((prow, fold, Wv, kc))
                for py in range(prow):
                    y = y0 + py0 + py
                    if not 0 <= y < H:
                        continue
                    for f in range(fold):
                        if n0 + f >= x.shape[0]:
                            continue
                        for px in range(Wv):
                            xx = px + dx
                            if 0 <= xx < W:
                                patch[py, f, px, :] = x[n0 + f, y, xx, c0:c0 + kc]
                flat = patch.reshape(prow * Wt, kc)
                for j in range(k):
                    for m in range(macc):
                        a = flat[j * Wt + m * 128: j * Wt + m * 128 + 128]
                        assert a.shape[0] == 128
                        acc[m * 128:(m + 1) * 128] += a @ wts[item].T
                    item += 1
            assert item == i.patch_items
            acc += bias
            for pix in range(macc * 128):
                y, xt = y0 + pix // Wt, pix % Wt
                f, xx = xt // Wv, xt % Wv
                if y < H and xx < W and n0 + f < n_valid:
                    if ph < 0:
                        out[n0 + f, y, xx] = acc[pix]
                    elif ph < 4:
                        out[n0 + f, 2 * y + (ph >> 1), 2 * xx + (ph & 1)] = acc[pix, :cout]
                    else:
                        out[n0 + f, 2 * y + ph - 4, 2 * xx] = acc[pix, :cout]
                        out[n0 + f, 2 * y + ph - 4, 2 * xx + 1] = acc[pix, cout:2 * cout]
    return out
