"""An oracle pin that does not depend on the oracle's own code (Next round 1c).

One XPBD substep of the reference's default cube WITH ground contacts — solver::step (src/solver.rs:3-17):
Rigid::integrate (src/rigid.rs:82-99), collision::ground (src/collision.rs:13-35, position-level friction through
Frame::delta, src/frame.rs:40-44), solver::solve (src/solver.rs:19-27) with Constraint::inverse_resitance / act
(src/constraint.rs:25-37) and Rigid::apply_impulse (src/rigid.rs:113-123), Rigid::derive (src/rigid.rs:101-109) —
derived here from the mathematics of those lines in 60-digit arithmetic (mpmath), with rotations written as rotation
MATRICES of unit quaternions rather than cgmath's `q * v` operator sequence.  The f64 oracle must reproduce it to
1e-12: that checks the oracle's formulas, operand order of the Gauss-Seidel sweep, the stale contact points and the
`(I^-1 r) x p` quirk against an independent statement of the reference, at a precision where no formula error hides.
(What it cannot check is cgmath's rounding order — only the reference binary could, and it cannot be built here.)"""
import mpmath as mp
import numpy as np

import oracle_lib as xo

mp.mp.dps = 60


def V(x):
    return mp.matrix([mp.mpf(float(t)) for t in x])


def rot_matrix(q):
    """Rotation matrix of the quaternion (x, y, z, s) / |q|."""
    x, y, z, s = q
    n = mp.sqrt(x * x + y * y + z * z + s * s)
    x, y, z, s = x / n, y / n, z / n, s / n
    return mp.matrix([[1 - 2 * (y * y + z * z), 2 * (x * y - s * z), 2 * (x * z + s * y)],
                      [2 * (x * y + s * z), 1 - 2 * (x * x + z * z), 2 * (y * z - s * x)],
                      [2 * (x * z - s * y), 2 * (y * z + s * x), 1 - 2 * (x * x + y * y)]])


def cross(a, b):
    return mp.matrix([a[1] * b[2] - a[2] * b[1], a[2] * b[0] - a[0] * b[2], a[0] * b[1] - a[1] * b[0]])


def dot(a, b):
    return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]


def hamilton(a, b):
    """Quaternion product, (x, y, z, s) layout."""
    ax, ay, az, as_ = a
    bx, by, bz, bs = b
    return (as_ * bx + ax * bs + ay * bz - az * by, as_ * by + ay * bs + az * bx - ax * bz,
            as_ * bz + az * bs + ax * by - ay * bx, as_ * bs - ax * bx - ay * by - az * bz)


def qnormalize(q):
    n = mp.sqrt(sum(t * t for t in q))
    return tuple(t / n for t in q)


class Body:
    def __init__(self, r38):
        r = [mp.mpf(float(t)) for t in r38]
        self.inv_m = r[0]
        self.Iinv = mp.matrix(3, 3)
        for c in range(3):
            for row in range(3):
                self.Iinv[row, c] = r[1 + 3 * c + row]  # column-major
        self.F_ext, self.F_int = mp.matrix(r[10:13]), mp.matrix(r[13:16])
        self.T_ext, self.T_int = mp.matrix(r[16:19]), mp.matrix(r[19:22])
        self.vel, self.ang = mp.matrix(r[22:25]), mp.matrix(r[25:28])
        self.com, self.pos = mp.matrix(r[28:31]), mp.matrix(r[31:34])
        self.rot = tuple(r[34:38])

    def frame(self):  # src/rigid.rs:75-80
        R = rot_matrix(self.rot)
        return self.pos + self.com - R * self.com, R

    def substep(self, verts, dt):
        past_pos, past_rot = self.pos.copy(), self.rot
        past_p, past_R = self.frame()
        R = rot_matrix(self.rot)
        # integrate, src/rigid.rs:82-99
        self.vel = self.vel + dt * (self.F_ext + R * self.F_int) * self.inv_m
        self.pos = self.pos + dt * self.vel
        self.ang = self.ang + dt * (self.Iinv * (self.T_ext + R * self.T_int))
        w = (self.ang[0], self.ang[1], self.ang[2], mp.mpf(0))
        dq = tuple(dt * mp.mpf("0.5") * t for t in hamilton(w, self.rot))
        self.rot = qnormalize(tuple(a + b for a, b in zip(self.rot, dq)))
        # ground, src/collision.rs:13-35 (constraints generated once, before the sweep)
        fp, fR = self.frame()
        cons = []
        for v in verts:
            p = fR * v + fp
            if p[2] >= 0:
                continue
            target = mp.matrix([p[0], p[1], 0])
            corr = target - p
            local = fR.T * (p - fp)            # Frame::inverse * global
            delta = p - (past_R * local + past_p)  # Frame::delta
            delta_t = delta - corr * (dot(delta, corr) / dot(corr, corr))
            cons.append((p, target - delta_t))
        # solve, src/solver.rs:19-27
        compliance = mp.mpf(1e-6) / (dt * dt)
        for c0, c1 in cons:
            diff = c1 - c0
            dist = mp.sqrt(dot(diff, diff))
            direction = diff / dist
            Rn = rot_matrix(self.rot)
            a = Rn.T * cross(c0 - (self.pos + self.com), direction)  # rotation.conjugate() * (r x n)
            wgt = self.inv_m + dot(self.Iinv * a, a)
            lam = (dist - 0) / (wgt + compliance)
            impulse = lam * direction
            # apply_impulse, src/rigid.rs:113-123: position first, then (I^-1 r) x p with the UPDATED position
            self.pos = self.pos + impulse * self.inv_m
            t = cross(self.Iinv * (c0 - (self.pos + self.com)), impulse)
            dq = tuple(mp.mpf("0.5") * x for x in hamilton((t[0], t[1], t[2], mp.mpf(0)), self.rot))
            self.rot = qnormalize(tuple(x + y for x, y in zip(self.rot, dq)))
        # derive, src/rigid.rs:101-109
        self.vel = (self.pos - past_pos) / dt
        conj = (-past_rot[0], -past_rot[1], -past_rot[2], past_rot[3])
        d = hamilton(self.rot, conj)
        if d[3] < 0:
            d = tuple(-x for x in d)
        self.ang = mp.matrix([2 * d[0] / dt, 2 * d[1] / dt, 2 * d[2] / dt])
        return len(cons)

    def dynamic(self):
        return np.array([float(x) for x in list(self.vel) + list(self.ang) + list(self.pos) + list(self.rot)])


def _start_state():
    """World::new's cube (src/world.rs:13,16-20) moved down to the ground, slightly tilted, yawed and moving: three corners
    of its bottom face are below z = 0 after the first integrate (a 3-constraint Gauss-Seidel sweep with friction offsets),
    one in the second substep, none afterwards."""
    a = xo.world_new()[0].copy()
    q = np.array([0.01, -0.004, 0.3, 1.0]); q /= np.linalg.norm(q)
    a[34:38] = q
    a[31:34] = [0.3, -0.2, -0.002]
    a[22:25] = [0.4, -0.25, -0.5]
    a[25:28] = [-0.2, 0.1, 0.5]
    return a


def test_one_substep_with_ground_contacts_matches_the_independent_derivation():
    cube = xo.builtin_hull(1)[0]
    verts = [V(v) for v in cube]
    a = _start_state()
    body = Body(a)
    dt64 = (1.0 / 60.0) / 25.0  # solver::step: dt / substep_count as f64
    dt = mp.mpf(dt64)
    r = a.reshape(1, 38).copy()
    contacts = []
    for k in range(1, 5):
        ncons = body.substep(verts, dt)
        # the oracle: ONE substep of dt64 (step_bodies divides frame dt by the substep count: 1 substep of dt64)
        n_oracle = xo.step_bodies(r, cube, dt64, 1, 1, threads=1)
        assert n_oracle == ncons, (k, n_oracle, ncons)
        contacts.append(ncons)
        want = body.dynamic()
        got = np.concatenate([r[0, 22:28], r[0, 31:38]])
        err = np.abs(got - want).max()
        assert err < 1e-12, (k, err, got - want)
        assert np.array_equal(r[0, :22], a[:22]) and np.array_equal(r[0, 28:31], a[28:31])  # constants untouched
    assert contacts == [3, 1, 0, 0]  # a multi-constraint sweep, a single constraint, free flight
