"""np_oracle.py — a second, independent CPU restatement of ray_marching.wgsl in vectorised numpy
float32 (TEST INFRASTRUCTURE ONLY; same rules as rm_oracle.c: nothing under ray-marching_b200/ may
import it).  It exists to cross-check oracle/rm_oracle.c bit for bit: numpy float32 arithmetic rounds
every add/mul/div/sqrt once and never fuses, i.e. the same IEEE semantics the C oracle is built
with (-ffp-contract=off).  PARITY UNPINNED by the reference itself (no tests/goldens upstream).

Follows /root/reference/src/ray_marching/ray_marching.wgsl: fs_main :36-76, ray_march :87-131,
calculate_normal :135-144, map_scene/eval_cmd_* :187-252 (+ extension opcodes 102/103, and the N4
extension: Plane = 2, Translation/Rotation/Scale Push/Pop = 200..205 on a position stack, at the opcode
values the reference's enum reserves in comments, builder.rs:8,16-23).
"""
from __future__ import annotations

import numpy as np

F = np.float32


def _f(x):
    return np.asarray(x, dtype=F)


def _min(a, b):  # WGSL wording: (b < a) ? b : a
    return np.where(b < a, b, a)


def _max(a, b):  # (a < b) ? b : a
    return np.where(a < b, b, a)


def map_scene(cmd_count, words, px, py, pz, max_dist=F(100.0)):
    """wgsl:187-252 over arrays of points."""
    if cmd_count == 0:
        return np.full(px.shape, max_dist, F)
    w = np.asarray(words, np.uint32)
    fb = w.view(F)
    stack = []
    pos_stack = []  # N4: saved evaluation points
    ptr = 0
    zero = F(0.0)
    for _ in range(cmd_count):
        t = int(w[ptr]); ptr += 1
        if t == 200:  # TranslationPush: pos = pos - t
            tx, ty, tz = fb[ptr:ptr + 3]; ptr += 3
            pos_stack.append((px, py, pz))
            px, py, pz = (px - tx).astype(F), (py - ty).astype(F), (pz - tz).astype(F)
            continue
        if t == 202:  # RotationPush (row-major R, object -> world): pos = R^T pos
            m = fb[ptr:ptr + 9]; ptr += 9
            pos_stack.append((px, py, pz))
            x = ((m[0] * px + m[3] * py) + m[6] * pz).astype(F)
            y = ((m[1] * px + m[4] * py) + m[7] * pz).astype(F)
            z = ((m[2] * px + m[5] * py) + m[8] * pz).astype(F)
            px, py, pz = x, y, z
            continue
        if t == 204:  # ScalePush: pos = pos / s
            sc = fb[ptr]; ptr += 1
            pos_stack.append((px, py, pz))
            px, py, pz = (px / sc).astype(F), (py / sc).astype(F), (pz / sc).astype(F)
            continue
        if t in (201, 203):  # Translation / Rotation Pop
            px, py, pz = pos_stack.pop()
            continue
        if t == 205:  # ScalePop: the child's distance was measured in scaled units
            sc = fb[ptr]; ptr += 1
            px, py, pz = pos_stack.pop()
            stack[-1] = (stack[-1] * sc).astype(F)
            continue
        if t == 2:  # plane: dot(pos, n) + h
            nx, ny, nz, h = fb[ptr:ptr + 4]; ptr += 4
            v = ((px * nx + py * ny) + pz * nz) + h
        elif t == 0:  # sphere :229-233
            cx, cy, cz, r = fb[ptr:ptr + 4]; ptr += 4
            dx, dy, dz = px - cx, py - cy, pz - cz
            v = np.sqrt((dx * dx + dy * dy) + dz * dz) - r
        elif t == 1:  # box :235-240
            cx, cy, cz, rx, ry, rz = fb[ptr:ptr + 6]; ptr += 6
            qx, qy, qz = np.abs(px - cx) - rx, np.abs(py - cy) - ry, np.abs(pz - cz) - rz
            mx, my, mz = _max(qx, zero), _max(qy, zero), _max(qz, zero)
            v = np.sqrt((mx * mx + my * my) + mz * mz) + _min(_max(qx, _max(qy, qz)), zero)
        elif t in (100, 101, 102):
            b = stack.pop(); a = stack.pop()
            v = _min(a, b) if t == 100 else (_max(a, -b) if t == 101 else _max(a, b))
        elif t == 103:
            k = fb[ptr]; ptr += 1
            b = stack.pop(); a = stack.pop()
            hh = _max(k - np.abs(a - b), zero) / k
            v = _min(a, b) - ((hh * hh) * k) * F(0.25)
        else:
            raise ValueError(f"bad opcode {t}")
        stack.append(v.astype(F))
    return stack.pop()


def _mat_vec(m, x, y, z, w):
    """column-major 4x4 times vec4 as a left-to-right sum of scaled columns."""
    out = []
    for r in range(4):
        acc = m[r] * x
        acc = acc + m[4 + r] * y
        acc = acc + m[8 + r] * z
        acc = acc + m[12 + r] * w
        out.append(acc.astype(F))
    return out


def render(uniforms: bytes, cmd_count: int, words, W: int, H: int, limits=(0.01, 100.0, 100)):
    """-> dict(rgba[H,W,4] u8, cls[H,W,16] u8, t[H,W,16] f32, counters[4])."""
    u = np.frombuffer(uniforms, F)
    ext, inv_proj, inv_view = u[0:2], u[4:20], u[20:36]
    min_dist, max_dist, max_iter = F(limits[0]), F(limits[1]), int(limits[2])
    ys, xs = np.meshgrid(np.arange(H, dtype=F), np.arange(W, dtype=F), indexing="ij")
    sx = ((xs + F(0.5)) / F(W) * F(2.0) - F(1.0)).astype(F)
    sy = (F(1.0) - (ys + F(0.5)) / F(H) * F(2.0)).astype(F)
    one, zero = np.ones_like(sx), np.zeros_like(sx)
    ro = _mat_vec(inv_view, zero, zero, zero, one)
    total = [np.zeros((H, W), F) for _ in range(3)]
    cls_out = np.zeros((H, W, 16), np.uint8)
    t_out = np.zeros((H, W, 16), F)
    n_step = n_hit = n_floor = 0
    with np.errstate(all="ignore"):
        for i in range(4):
            for j in range(4):
                offx = (F(i) + F(0.5)) / F(4.0) - F(0.5)
                offy = (F(j) + F(0.5)) / F(4.0) - F(0.5)
                ax = F(offx / ext[0]) * F(2.0)
                ay = F(offy / ext[1]) * F(2.0)
                pv = _mat_vec(inv_proj, (sx + ax).astype(F), (sy + ay).astype(F), -one, one)
                pw = _mat_vec(inv_view, *pv)
                d4 = [(pw[k] - ro[k]).astype(F) for k in range(4)]
                ln = np.sqrt(((d4[0] * d4[0] + d4[1] * d4[1]) + d4[2] * d4[2]) + d4[3] * d4[3]).astype(F)
                d = [(d4[k] / ln).astype(F) for k in range(3)]
                dist = np.zeros((H, W), F)
                active = np.ones((H, W), bool)
                hit = np.zeros((H, W), bool)
                col = [np.zeros((H, W), F) for _ in range(3)]
                for _ in range(max_iter):
                    if not active.any():
                        break
                    px, py, pz = [(ro[k] + d[k] * dist).astype(F) for k in range(3)]
                    s = map_scene(cmd_count, words, px, py, pz, max_dist)
                    n_step += int(active.sum())
                    h_now = active & (s < min_dist)
                    if h_now.any():
                        e = F(0.0001)
                        K = [(1, -1, -1), (-1, -1, 1), (-1, 1, -1), (1, 1, 1)]
                        n = None
                        for kx, ky, kz in K:
                            m = map_scene(cmd_count, words, (px + F(kx) * e).astype(F), (py + F(ky) * e).astype(F),
                                          (pz + F(kz) * e).astype(F), max_dist)
                            term = [F(kx) * m, F(ky) * m, F(kz) * m]
                            n = term if n is None else [(n[q] + term[q]).astype(F) for q in range(3)]
                        nl = np.sqrt((n[0] * n[0] + n[1] * n[1]) + n[2] * n[2]).astype(F)
                        n = [(n[q] / nl).astype(F) for q in range(3)]
                        l = [(px - F(2.0)).astype(F), (py - F(-5.0)).astype(F), (pz - F(3.0)).astype(F)]
                        ll = np.sqrt((l[0] * l[0] + l[1] * l[1]) + l[2] * l[2]).astype(F)
                        l = [(l[q] / ll).astype(F) for q in range(3)]
                        diff = _max(F(0.02), ((n[0] * l[0] + n[1] * l[1]) + n[2] * l[2]).astype(F))
                        for q, a in enumerate((F(0.4), F(0.7), F(0.1))):
                            col[q] = np.where(h_now, (a * diff).astype(F), col[q])
                        hit |= h_now
                        n_hit += int(h_now.sum())
                    active &= ~h_now
                    brk = active & (s > max_dist)
                    active &= ~brk
                    dist = np.where(active, (dist + s).astype(F), dist)
                fd = ((F(-1.5) - ro[1]) / d[1]).astype(F)
                floor = (~hit) & (fd > 0)
                fx = (ro[0] + d[0] * fd).astype(F)
                fz = (ro[2] + d[2] * fd).astype(F)

                def sat(v):
                    r = np.rint(v)
                    r = np.where(np.isnan(r), 0, r)
                    return np.clip(r, -2147483648.0, 2147483647.0).astype(np.int64).astype(np.int32)
                c = ((sat((fx + F(0.5)).astype(F)) ^ sat((fz + F(0.5)).astype(F))) & 1).astype(F)
                fcol = [(F(0.1) + F(0.2) * c).astype(F), (F(0.1) + F(0.2) * c).astype(F), (F(0.2) + F(0.2) * c).astype(F)]
                n_floor += int(floor.sum())
                for q in range(3):
                    cq = np.where(hit, col[q], np.where(floor, fcol[q], F(0.0))).astype(F)
                    total[q] = (total[q] + np.sqrt(cq)).astype(F)
                cls_out[:, :, i * 4 + j] = np.where(hit, 2, np.where(floor, 1, 0))
                t_out[:, :, i * 4 + j] = np.where(hit, dist, fd)
    rgba = np.zeros((H, W, 4), np.uint8)
    for q in range(3):
        v = (total[q] / F(16.0)).astype(F)
        v = np.where(np.isnan(v), F(0), np.clip(v, F(0), F(1)))
        rgba[:, :, q] = np.floor(v * F(255.0) + F(0.5)).astype(np.uint8)
    rgba[:, :, 3] = 255
    return {"rgba": rgba, "cls": cls_out, "t": t_out,
            "counters": np.array([n_step, n_hit, n_floor, W * H * 16], np.uint64)}
