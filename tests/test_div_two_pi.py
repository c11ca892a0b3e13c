"""The shader divides through double-precision reciprocals (mcpt_device.cuh: fdiv_two_pi, operator/(f4, float)); the claim is
that the narrowed products carry the bits of the IEEE float quotients the reference computes (shade.cl:120,143; the
normalisations of shade.cl:40-59).  tools/src/div_two_pi_check.c enumerates the division by 2 pi over float bit patterns
(all 2^32 with stride 1: 3 s on 16 cores; a stride of 61 here) and samples the variable-divisor case on random pairs."""
import json
import os
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_reciprocal_products_round_like_the_ieee_quotients(tmp_path):
    exe = str(tmp_path / "div_two_pi_check")
    subprocess.run(["gcc", "-O2", "-fopenmp", "-o", exe, os.path.join(ROOT, "tools", "src", "div_two_pi_check.c"), "-lm"], check=True)
    out = subprocess.run([exe, "61", "40000000"], stdout=subprocess.PIPE, text=True, check=True).stdout.strip().split("\n")
    pairs, single = json.loads(out[0]), json.loads(out[1])
    assert pairs["pairs_checked"] > 3.9e7 and pairs["pair_mismatches"] == 0
    assert single["checked"] > 7e7 and single["mismatches"] == 0
    # the constant in the device code is the reciprocal the check used
    src = open(os.path.join(ROOT, "montecarlopathtracing_b200", "csrc", "cuda", "mcpt_device.cuh")).read()
    assert single["reciprocal"] in src
