from lbm_ferrofluid_b200.simulation import SimulationParameters, SimulationRunner  # noqa: F401
