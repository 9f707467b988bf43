"""The tuned kernels divide by h*h with a corrected reciprocal when h*h is not a power of two (gs_device.cuh
div_hh<AR_RCP>: q = s*y, two fma corrections).  Markstein's theorem makes the result the correctly rounded quotient;
this test checks the exact sequence against the IEEE divide on the CPU for every h = 1/(n+1), n <= 4096."""
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))


def test_corrected_reciprocal_equals_ieee_divide(tmp_path):
    exe = str(tmp_path / "rcp_check")
    subprocess.run(["gcc", "-O2", "-ffp-contract=off", os.path.join(HERE, "rcp_division_check.c"), "-lm", "-o", exe], check=True)
    r = subprocess.run([exe, "4096", "20000"], capture_output=True, text=True, timeout=300)
    bad, total = map(int, r.stdout.split())
    assert total == 4096 * 20000
    assert bad == 0 and r.returncode == 0, f"{bad} of {total} quotients differ from the fp64 divide"
