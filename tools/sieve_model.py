"""numpy model of the sieve's cheap, certified rejection (csrc/update_depth.cu, phase A0).

The CUDA sieve decides most pixels without the reference's exact arithmetic: it evaluates the five
key samples of epipolar.rs:22-36 / interpolation.rs:9-43 in f32 on an f32 copy of the key image,
bounds the distance to the reference's f64 result and finishes a pixel only when the bound puts it
on one side of the test for certain.  This file restates that decision in numpy (float32 where the
kernel uses f32) so that the bound can be checked against the oracle on the CPU:

    certain(-6)  must imply  oracle flag == -6       certain(-2)  must imply  oracle flag == -2

    python tools/sieve_model.py            # certification rates on a few synthetic keyframes
"""
import numpy as np

EPS_INV = np.finfo(np.float64).eps
H_MAX = 256.0          # longest half segment (pixels) the cheap path accepts


def cheap_sieve(key_img, key_cam, T_wk, ref_cams, T_wr, age, prior_d, prior_v, params):
    """One keyframe.  Returns an int8 map: 0 = undecided (exact path), -6 / -2 / -9 / -7 / -1 / -8 =
    decided here.  (-9, -7, -1, -8 are decided exactly, as in the kernel.)"""
    f32 = np.float32
    H, W = key_img.shape
    R = T_wr.shape[0]
    vmin, vmax = 1. / (params[1] + EPS_INV), 1. / (params[0] + EPS_INV)
    step, min_grad = params[4], params[5]
    fx, fy, ox, oy = key_cam
    out = np.zeros((H, W), dtype=np.int8)
    with np.errstate(all="ignore"):
        img32 = key_img.astype(f32)
        amax = np.abs(key_img).max()
        if not np.isfinite(amax) or amax > 1e30:
            amax = np.inf
        ys, xs = np.mgrid[0:H, 0:W]
        xk = (xs - ox) / fx
        yk = (ys - oy) / fy
        a = age.astype(np.int64)
        out[a == 0] = -9
        live = (a > 0) & (a <= R)
        idp = 1. / (prior_d + EPS_INV)
        neg = live & (idp <= 0)
        out[neg] = -7
        lo, hi = idp - 2 * prior_v, idp + 2 * prior_v
        oor = live & ~neg & ((hi <= vmin) | (vmax <= lo))
        out[oor] = -1
        live &= ~neg & ~oor
        kd = 1. / (idp + EPS_INV)
        r = np.clip(R - a, 0, R - 1)
        Trk = np.stack([np.linalg.inv(T_wr[i]) @ T_wk for i in range(R)])   # model only: the kernel has the exact LU
        t_wr = T_wr[:, :3, 3]
        ek = (T_wk[:3, :3].T @ (t_wr - T_wk[:3, 3]).T).T
        ex, ey = ek[:, 0] / (ek[:, 2] + 1e-16), ek[:, 1] / (ek[:, 2] + 1e-16)
        T8, T9, T10, T11 = (Trk[:, 2, j][r] for j in range(4))
        rz = T8 * (xk * kd) + T9 * (yk * kd) + T10 * (1. * kd) + T11
        nrd = live & (rz <= 0)
        out[nrd] = -8
        live &= ~nrd
        # ---- approximate part
        kstep = idp * (rz + EPS_INV) * step
        kdx, kdy = xk - ex[r], yk - ey[r]
        n2 = kdx * kdx + kdy * kdy
        rn = 1. / np.sqrt(n2)
        ax = (kstep * (kdx * rn) * fx).astype(f32)
        ay = (kstep * (kdy * rn) * fy).astype(f32)
        hext = 2 * np.maximum(np.abs(ax), np.abs(ay))
        sane = live & (hext < f32(H_MAX))
        xf, yf = xs.astype(f32), ys.astype(f32)
        cx = [f32(k - 2) * ax + xf for k in range(5)]     # fmaf in the kernel; numpy rounds the product first --
        cy = [f32(k - 2) * ay + yf for k in range(5)]     # the model's own bound below covers either
        # coordinate error bound (pixels): f32 rounding of a and of the sum, the reference's f64 rounding
        delta = (H_MAX + max(W, H)) * 2.0 ** -22
        dm = f32(max(delta, 1e-3))
        wm, hm = f32(W - 1), f32(H - 1)

        def inside(x, y):
            return (x >= dm) & (x <= wm - dm) & (y >= dm) & (y <= hm - dm)

        def outside(x, y):
            return (x < -dm) | (x > wm + dm) | (y < -dm) | (y > hm + dm)

        sure_in = sane & inside(cx[0], cy[0]) & inside(cx[4], cy[4])
        sure_out = sane & (outside(cx[0], cy[0]) | outside(cx[4], cy[4]))
        out[sure_out] = -2
        # ---- five samples in f32
        def sample(x, y):
            x = np.where(sure_in, x, f32(1)); y = np.where(sure_in, y, f32(1))
            lx, ly = np.floor(x), np.floor(y)
            wx, wy = x - lx, y - ly
            ix, iy = lx.astype(np.int64), ly.astype(np.int64)
            a_, b_ = img32[iy, ix], img32[iy, ix + 1]
            c_, d_ = img32[iy + 1, ix], img32[iy + 1, ix + 1]
            top = a_ + wx * (b_ - a_)
            bot = c_ + wx * (d_ - c_)
            return (top + wy * (bot - top)).astype(f32)

        s = [sample(cx[k], cy[k]) if k != 2 else img32 for k in range(5)]
        S = f32(0)
        for k in range(4):
            d = s[k + 1] - s[k]
            S = S + d * d
        # |sample_f32 - sample_ref| <= eps_s ; per difference e = 2 eps_s
        eps_s = 4.0 * delta * amax + 16 * 2.0 ** -24 * amax
        e = 2.0 * eps_s
        thr = min_grad * (1 - 1e-6) - 2.0 * e
        thr2 = f32(thr * thr * (1 - 1e-6)) if thr > 0 else f32(-1)
        sure_low = sure_in & (S < thr2)
        out[sure_low] = -6
    return out


def main():
    import sys, os
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    import oracle as ora
    from tadataka_b200 import scene
    cases = [("C2 seed 2000", scene.make_keyframe(2000, R=8)),
             ("C2 seed 2003", scene.make_keyframe(2003, R=8)),
             ("C1 seed 1000", scene.make_pair(1000)),
             ("C4 720p set T", scene.make_pair(4000, H=720, W=1280, cam=scene.CAM_720P, pset=scene.SET_T))]
    for name, it in cases:
        b = scene.stack_keyframes([it])
        _, _, f = ora.update_depth_batch(b["key_img"], b["key_cam"], b["T_wk"], b["ref_imgs"], b["ref_cams"],
                                         b["T_wr"], b["age"], b["prior_depth"], b["prior_variance"],
                                         ora.make_params(*b["params"]))[:3]
        f = f[0]
        m = cheap_sieve(it["key_img"], it["key_cam"], it["T_wk"], it["ref_cams"], it["T_wr"], it["age"],
                        it["prior_depth"], it["prior_variance"], it["params"])
        dec = m != 0
        wrong = dec & (m != f)
        print(f"{name}: decided {dec.mean():.4f} of the pixels; undecided {int((~dec).sum())} "
              f"(oracle candidates {(f >= -5).sum() - (f == -2).sum() - (f == -1).sum()}, oracle -6: {(f == -6).sum()}, "
              f"-6 left undecided: {((f == -6) & ~dec).sum()}); contradictions: {int(wrong.sum())}")


if __name__ == "__main__":
    main()
