"""The reference driver's time loop (driver/pace/driver/driver.py:639-662) on pace_b200, with the state on the device:

    for step in range(n):
        dycore.step_dynamics(state)                      -> engine.step_dynamics()
        dycore_to_physics(dycore_state, physics_state)   -> engine.dycore_to_physics()
        physics(physics_state)                           -> `physics(engine)`: the caller's, reads / writes "phy.*"
        end_of_step_update(dycore_state, phy_state, ...) -> engine.update_atmos_state()
        diagnostics.store(time, state)                   -> diagnostics.store(time)

`physics` below is a stand-in (the parameterisations are not part of this library): it returns the state it was
handed with a small relaxation of the temperature, which is enough to exercise every branch of the coupling.
"""
import torch


def toy_physics(engine, dt):
    """Fills phy.prsi and phy.physics_updated_* from the copied state, on the device."""
    t = engine.core.t
    nz = engine.npz
    ptop = float(engine.cfg["ptop"])
    t["phy.prsi"][:, 0] = ptop
    t["phy.prsi"][:, 1:] = ptop + torch.cumsum(t["phy.delp"][:, :nz], dim=1)
    for name in ("ua", "va", "qliquid", "qrain", "qsnow", "qice", "qgraupel"):
        t["phy.physics_updated_" + name].copy_(t["phy." + name])
    t["phy.physics_updated_specific_humidity"].copy_(t["phy.qvapor"])
    # Newtonian relaxation of the temperature towards 250 K with a 10-day time scale
    t["phy.physics_updated_pt"].copy_(t["phy.pt"] + (250.0 - t["phy.pt"]) * (dt / 864000.0))


def run(engine, driver_grids, n_steps, diagnostics=None, physics=toy_physics):
    dt = float(engine.cfg["dt_atmos"])
    engine.setup_coupling(driver_grids)
    for step in range(n_steps):
        engine.step_dynamics()
        engine.dycore_to_physics()
        physics(engine, dt)
        engine.update_atmos_state()
        if diagnostics is not None:
            diagnostics.store((step + 1) * dt)
    engine.core.sync()
