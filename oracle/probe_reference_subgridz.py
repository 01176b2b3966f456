"""TEST INFRASTRUCTURE ONLY — can the UNMODIFIED reference run DycoreToPhysics with do_dry_convective_adjust=True?

    python oracle/probe_reference_subgridz.py

Builds the reference's DycoreToPhysics the way its driver does (driver/pace/driver/driver.py:519-525) with the dry
convective adjustment switched on (fv_sg_adj = 600, n_sponge = 48, the values of
driver/examples/configs/tropical_read_restart_fortran.yml) and calls it on the DycoreState after one dycore step, C12,
one rank.  Outcome in this container (profiles/r02_reference_subgridz_probe.txt): the call raises
`TypeError: 'Quantity' object is not subscriptable` at pyFV3/stencils/fv_subgridz.py:874 (`state.pe[self._is, self._js,
0]`: DycoreState holds Quantities, which cannot be indexed).  The driver path of this option is therefore not
reachable in the reference either; pace_b200.dropin.DycoreToPhysics refuses it with NotImplementedError.
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from oracle.refshim.bootstrap import bootstrap  # noqa: E402

bootstrap()


def main():
    import dataclasses
    import traceback

    from ndsl.comm.null_comm import NullComm
    from ndsl.constants import X_DIM, Y_DIM, Z_DIM
    from pySHiELD.update import update_atmos_state

    from oracle.refshim.run_reference import build_dycore, default_config_kwargs

    dycore, state, grid_data, damping, metric_terms, config, communicator = build_dycore(
        NullComm(rank=0, total_ranks=6, fill_value=0.0), default_config_kwargs(npx=13, n_split=1, fv_sg_adj=600, n_sponge=48))
    dycore.step_dynamics(state)
    qf, sf = build_dycore.extras["quantity_factory"], build_dycore.extras["stencil_factory"]

    @dataclasses.dataclass
    class Tend:
        u_dt: object
        v_dt: object
        pt_dt: object

    tend = Tend(*(qf.zeros([X_DIM, Y_DIM, Z_DIM], "unknown") for _ in range(3)))
    d2p = update_atmos_state.DycoreToPhysics(stencil_factory=sf, quantity_factory=qf, dycore_config=config,
                                             do_dry_convective_adjust=True, dycore_only=True)
    try:
        d2p(dycore_state=state, physics_state=None, tendency_state=tend, timestep=float(config.dt_atmos))
        print("DycoreToPhysics(do_dry_convective_adjust=True) ran")
    except Exception as e:  # noqa: BLE001
        tb = traceback.extract_tb(e.__traceback__)
        where = [f for f in tb if "/root/reference/" in f.filename][-1]
        print("DycoreToPhysics(do_dry_convective_adjust=True) raised %s: %s" % (type(e).__name__, e))
        print("at %s:%d  %s" % (where.filename.replace("/root/reference/", ""), where.lineno, where.line))


if __name__ == "__main__":
    main()
