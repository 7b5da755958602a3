"""Parameter bag and operator factory (reference: src/LBM/simulation/simulation_{parameters,runner}.py)."""
from __future__ import annotations

import math
from typing import List

import torch

from .collision import (LBMCollision2d, LBMCollision3d, LBMCollisionHCZ2d, LBMCollisionHCZ3d, LBMCollisionMRT2d,
                        LBMCollisionSC2d)
from .macro_compute import LBMMacroCompute2d, LBMMacroCompute3d
from .magnetic import LBMMagnetic2d, LBMMagnetic3d
from .propagation import LBMPropagation2d, LBMPropagation3d


class SimulationParameters(object):
    """Same 16 keyword arguments, attributes and helpers as simulation_parameters.py:6-83."""

    def __init__(self, dtype=torch.float32, dim: int = 2, dt: float = 1.0, device=torch.device("cpu"),
                 simulation_size: List[int] = [1, 1, 256, 256], density_gas: float = 0.0,
                 density_fluid: float = 1.0, contact_angle=0.75 * math.pi, Q: float = 9, tau: float = 1.0,
                 rho_gas: float = 0.038, rho_fluid: float = 0.265, kappa: float = 0.08, tau_g: float = 0.7,
                 tau_f: float = 0.7, k: float = 1.0):
        if dim == 2 and len(simulation_size) != 4:
            raise ValueError("For 2d simulation simulation size should have 4 parameters B x C x H x W")
        self.dim, self.dtype, self.dt, self.device = dim, dtype, dt, device
        self.frame, self.time_per_frame, self.frame_length, self.time_total = 0, 0, 1.0, 0
        self.simulation_size = simulation_size
        self.density_gas, self.density_fluid, self.contact_angle = density_gas, density_fluid, contact_angle
        self.Q, self.tau = Q, tau
        self.rho_gas, self.rho_fluid = rho_gas, rho_fluid
        self.kappa, self.tau_g, self.tau_f = kappa, tau_g, tau_f
        self.k = k  # susceptibility; like the reference, never read by an op (0.33 is hard-coded there)

    def step(self):
        """Advance the host-side clocks by one time step (simulation_parameters.py:57-67)."""
        self.time_per_frame += self.dt
        self.time_total += self.dt
        if self.time_per_frame >= self.frame_length:
            self.frame += 1
            self.time_total = self.frame * self.frame_length
            self.time_per_frame = 0

    def get_dx(self):
        return 1.0 / max(self.simulation_size)

    def is_2d(self):
        return self.dim == 2

    def is_3d(self):
        return self.dim == 3

    def set_device(self, device: str = "cuda"):
        if device not in ["cuda", "cpu"]:
            raise ValueError("Set_device: device {} must be either cuda or cpu.".format(device))
        self.device = torch.device(device)


class SimulationRunner(object):
    """Factory of the operator objects (simulation_runner.py:22-194).  3-D MRT and 3-D Shan-Chen return
    ``None`` exactly like the reference (simulation_runner.py:107-108, 126-127)."""

    def __init__(self, parameters: SimulationParameters):
        self.parameters = parameters

    def _kw(self, full: bool = True):
        p = self.parameters
        kw = dict(Q=p.Q, tau=p.tau, density_liquid=p.density_fluid, density_gas=p.density_gas,
                  rho_liquid=p.rho_fluid, rho_gas=p.rho_gas, dtype=p.dtype, device=p.device)
        if full:
            kw.update(kappa=p.kappa, tau_f=p.tau_f, tau_g=p.tau_g, contact_angle=p.contact_angle)
        return kw

    def _make(self, cls2d, cls3d, full=True):
        cls = cls2d if self.parameters.is_2d() else cls3d
        return cls(**self._kw(full)) if cls is not None else None

    def create_propagation(self):
        return LBMPropagation2d() if self.parameters.is_2d() else LBMPropagation3d()

    def create_macro_compute(self):
        return self._make(LBMMacroCompute2d, LBMMacroCompute3d, full=False)

    def create_collision(self):
        return self._make(LBMCollision2d, LBMCollision3d)

    def create_collision_MRT(self):
        return self._make(LBMCollisionMRT2d, None)

    def create_collision_SC(self):
        return self._make(LBMCollisionSC2d, None)

    def create_collision_HCZ(self):
        return self._make(LBMCollisionHCZ2d, LBMCollisionHCZ3d)

    def create_LBM_magnetic(self):
        return self._make(LBMMagnetic2d, LBMMagnetic3d)

    def step(self):
        self.parameters.step()
