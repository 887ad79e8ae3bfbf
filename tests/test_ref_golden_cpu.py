"""The CPU oracle against vectors produced by the REFERENCE's own Python statements of the path
(tests/golden/ref_golden.npz, generated in the build container by tests/golden/make_ref_golden.py from
/root/reference/unittest/python/crocoddyl and bindings/python/crocoddyl/utils). This is what pins the oracle."""
import pytest

from golden import ref_checks as rc


@pytest.mark.parametrize("cname", rc.names("unicycle_node") + rc.names("lqr_node"))
def test_oracle_node_calc_calcdiff_vs_reference(cname):
    rc.check_node(rc.OracleEngine, cname)


@pytest.mark.parametrize("cname", rc.names("step_"))
def test_oracle_direction_and_trystep_vs_reference(cname):
    rc.check_step(rc.OracleEngine, cname)


@pytest.mark.parametrize("cname", rc.names("kkt_"))
def test_oracle_ddp_reaches_reference_kkt_solution(cname):
    rc.check_kkt(rc.OracleEngine, cname)


@pytest.mark.parametrize("cname", rc.names("solve_"))
def test_oracle_solve_sequence_vs_reference(cname):
    rc.check_solve(rc.OracleEngine, cname)
