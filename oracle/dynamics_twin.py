"""TEST INFRASTRUCTURE ONLY — float64 CPU twin of the planar skid-steer dynamics.

The reference has NO dynamics code: physics is the closed-source PhysX articulation stepped at
src/my_envs/KAMMA_medium.py:646-653.  The B200 path replaces it with the explicit model below; this
file is the float64 statement of *those same equations* that the CUDA kernel
(options_in_simulation_b200/csrc/env_step.cu: ``skid_substep``) is checked against (<= 1e-6).
PhysX trajectory agreement is "unavailable on this box" (DESIGN.md §4).

Parameters (file:line in /root/reference):
  body 21 kg + 4 x 1 kg wheels          src/my_assets/urdf/origin_sym.urdf:11,34
  body Izz 1.01, wheel inertia 0.01     src/my_assets/urdf/origin_sym.urdf:12,35
  wheel radius 0.1175                   src/my_assets/urdf/origin_sym.urdf:40 ; KAMMA_medium.py:79
  wheel centres (+-0.205, +-0.25)       src/my_assets/urdf/origin_sym.urdf:115-139
  friction 0.85                         train/train_kamma_4d_nofifo_random_RLIL.py:66-67
  g 9.81, sim dt 0.01, decimation 10    src/my_envs/KAMMA_medium.py:84,82,75
  max wheel speed 64 rad/s              src/my_robots/origin_v18.py:70
  DOF order [fl, rl, fr, rr]            src/my_envs/KAMMA_medium.py:623

Model, per sub-step h (body-frame twist (vx, vy, r), wheel speeds w_i, pose (x, y, psi)):
  contact velocity of wheel i at (x_i, y_i):  u_i = vx - r*y_i ,  l_i = vy + r*x_i
  slip                                        sx_i = R*w_i - u_i ,  sy_i = -l_i
  regularised Coulomb gain (lagged)           k_i = mu*N_i / sqrt(sx_i^2 + sy_i^2 + eps^2),  N_i = m*g/4
  forces                                      Fx_i = k_i*(R*w_i' - u_i') ,  Fy_i = -k_i*l_i'
  wheels   Iw*(w_i' - w_i)/h = tau_i - R*Fx_i - c*w_i'
  body     m*(vx' - vx)/h =  m*r*vy + sum Fx_i
           m*(vy' - vy)/h = -m*r*vx + sum Fy_i
           J*(r'  - r )/h = sum (x_i*Fy_i - y_i*Fx_i)
  i.e. linearly-implicit Euler: the gains k_i and the Coriolis terms are evaluated at the old state,
  everything else at the new one.  Eliminating w_i' leaves a symmetric positive-definite 3x3 system
  in (vx', vy', r') that is solved in closed form.  Then
           w_i' = clamp(w_i + alpha_i + beta_i*(u_i' - u_i), +-w_max)
  (the code solves for the increments (vx'-vx, vy'-vy, r'-r), which is the same system with every
  right-hand-side term O(acceleration); that form keeps float32 within 1e-6 of float64)
           psi' = psi + h*r' ;  x' = x + h*(vx'*cos psi' - vy'*sin psi') ;  y' = y + h*(vx'*sin psi' + vy'*cos psi')
"""
from __future__ import annotations

from dataclasses import dataclass

import numpy as np


@dataclass(frozen=True)
class SkidSteerParams:
    mass: float = 25.0                 # 21 + 4*1
    yaw_inertia: float = 1.01 + 4 * 0.01 + 4 * 1.0 * (0.205**2 + 0.25**2)   # = 1.4681
    wheel_inertia: float = 0.01
    wheel_radius: float = 0.1175
    half_wheelbase: float = 0.205      # |x_i|
    half_track: float = 0.25           # |y_i|
    mu: float = 0.85
    gravity: float = 9.81
    slip_eps: float = 0.05             # m/s, Coulomb regularisation
    wheel_damping: float = 0.005       # N*m*s/rad ("control.damping", KAMMA_medium.py:77)
    wheel_speed_max: float = 64.0
    sim_dt: float = 0.01
    decimation: int = 10

    @property
    def wheel_x(self):
        return (self.half_wheelbase, -self.half_wheelbase, self.half_wheelbase, -self.half_wheelbase)

    @property
    def wheel_y(self):
        return (self.half_track, self.half_track, -self.half_track, -self.half_track)


def skid_substep(state, tau, p: SkidSteerParams, dtype=np.float64):
    """One sub-step.  ``state`` = dict of arrays px, py, yaw, vx, vy, wz, w (shape [N,4]); ``tau`` [N,4].
    Returns a new dict.  ``dtype`` float64 is the oracle; float32 mimics the kernel's arithmetic."""
    f = dtype
    h = f(p.sim_dt)
    m, J, Iw, R = f(p.mass), f(p.yaw_inertia), f(p.wheel_inertia), f(p.wheel_radius)
    c = f(p.wheel_damping)
    muN = f(p.mu * p.mass * p.gravity / 4.0)
    eps2 = f(p.slip_eps * p.slip_eps)
    vx, vy, r = state["vx"].astype(f), state["vy"].astype(f), state["wz"].astype(f)
    w = state["w"].astype(f)
    tau = tau.astype(f)

    m_h = m / h
    J_h = J / h
    Iw_h = Iw / h + c
    A00 = np.full_like(vx, m_h)
    A11 = np.full_like(vx, m_h)
    A22 = np.full_like(vx, J_h)
    A02 = np.zeros_like(vx)
    A12 = np.zeros_like(vx)
    # increment form: unknowns are (dvx, dvy, dr); every rhs term is O(acceleration), not O(velocity),
    # which keeps the float32 kernel within 1e-6 of this float64 statement
    b0 = m * (r * vy)
    b1 = -(m * (r * vx))
    b2 = np.zeros_like(vx)
    al_l, be_l = [], []
    for i in range(4):
        xi, yi = f(p.wheel_x[i]), f(p.wheel_y[i])
        u = vx - r * yi
        l = vy + r * xi
        sx = R * w[:, i] - u
        k = muN / np.sqrt(sx * sx + l * l + eps2)
        E = Iw_h + (R * R) * k
        Rk = R * k
        al = (tau[:, i] - c * w[:, i] - Rk * sx) / E
        be = Rk / E
        g = k * Iw_h / E
        phi = k * (sx + R * al)
        kl = k * l
        A00 = A00 + g
        A11 = A11 + k
        A22 = A22 + (xi * xi) * k + (yi * yi) * g
        A02 = A02 - yi * g
        A12 = A12 + xi * k
        b0 = b0 + phi
        b1 = b1 - kl
        b2 = b2 - xi * kl - yi * phi
        al_l.append(al)
        be_l.append(be)
    # symmetric 3x3 solve (A01 = 0): eliminate dvx, dvy into the yaw row
    i00 = f(1.0) / A00
    i11 = f(1.0) / A11
    s22 = A22 - A02 * A02 * i00 - A12 * A12 * i11
    dr = (b2 - A02 * b0 * i00 - A12 * b1 * i11) / s22
    dvx = (b0 - A02 * dr) * i00
    dvy = (b1 - A12 * dr) * i11
    vxn = vx + dvx
    vyn = vy + dvy
    rn = r + dr
    wn = np.empty_like(w)
    wmax = f(p.wheel_speed_max)
    for i in range(4):
        yi = f(p.wheel_y[i])
        du = dvx - dr * yi
        wn[:, i] = np.clip(w[:, i] + (al_l[i] + be_l[i] * du), -wmax, wmax)
    yaw = state["yaw"].astype(f) + h * rn
    cs, sn = np.cos(yaw), np.sin(yaw)
    px = state["px"].astype(f) + h * (vxn * cs - vyn * sn)
    py = state["py"].astype(f) + h * (vxn * sn + vyn * cs)
    return {"px": px, "py": py, "yaw": yaw, "vx": vxn, "vy": vyn, "wz": rn, "w": wn}


def skid_step(state, tau, p: SkidSteerParams = SkidSteerParams(), dtype=np.float64):
    """One policy step = ``decimation`` sub-steps with constant wheel torques (KAMMA_medium.py:646-653)."""
    s = {k: np.asarray(v) for k, v in state.items()}
    for _ in range(p.decimation):
        s = skid_substep(s, tau, p, dtype)
    return s


def random_state(n: int, rng: np.random.Generator):
    """Synthetic state per SURVEY.md §8(d)."""
    vx = rng.uniform(0.0, 2.0, n)
    return {
        "px": rng.uniform(-0.5, 0.5, n),
        "py": rng.uniform(-0.5, 0.5, n),
        "yaw": rng.uniform(0.0, 2 * np.pi, n),
        "vx": vx,
        "vy": rng.normal(0.0, 0.05, n),
        "wz": rng.normal(0.0, 0.05, n),
        "w": vx[:, None] / 0.1175 + rng.normal(0.0, 0.5, (n, 4)),
    }
