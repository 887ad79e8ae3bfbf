"""The CUDA path (through the C ABI) against vectors produced by the REFERENCE's own Python statements of the path
(tests/golden/ref_golden.npz; see tests/golden/make_ref_golden.py). Needs neither the oracle nor /root/reference:
the product is compared with the reference directly. Same checks and tolerances as the oracle's CPU test."""
import pytest

from golden import ref_checks as rc

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("cname", rc.names("unicycle_node") + rc.names("lqr_node"))
def test_node_calc_calcdiff_vs_reference(cname):
    rc.check_node(rc.GpuEngine, cname)


@pytest.mark.parametrize("cname", rc.names("step_"))
def test_direction_and_trystep_vs_reference(cname):
    rc.check_step(rc.GpuEngine, cname)


@pytest.mark.parametrize("cname", rc.names("kkt_"))
def test_ddp_reaches_reference_kkt_solution(cname):
    rc.check_kkt(rc.GpuEngine, cname)


@pytest.mark.parametrize("cname", rc.names("solve_"))
def test_solve_sequence_vs_reference(cname):
    rc.check_solve(rc.GpuEngine, cname)
