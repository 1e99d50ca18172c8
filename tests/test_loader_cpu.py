"""host/load.cc (the drop-in for the reference's src/load.cc) against the oracle's restatement of the same
parser, on the reference clouds and on the awkward lines sscanf("%lf,%lf,%lf") tolerates.  No GPU needed."""
import os
import subprocess

import numpy as np
import pytest

from conftest import PKG_DIR, cloud_text_path
from oracle import oracle as O

LOADCHECK = os.path.join(PKG_DIR, "bin", "loadcheck")

TRICKY = (
    "Points:0,Points:1,Points:2\n"
    "1.5,2.5,3.5\n"
    "-0.001,+4e-3,5E2\n"                 # '+' and exponents: general path
    "  7.25, 8.5,9\n"                     # leading blanks (sscanf skips them before %lf)
    "1,2\n"                               # short line: z stays 0
    "3\n"                                 # only x
    "\n"                                  # empty line: all zeros
    "0.1,0.2,0.3,99.0,12\n"               # extra columns (cow_tr2 has five)
    "1e-320,2,3\n"                        # denormal
    "abc,1,2\n"                           # not a number: zeros
    "4,x,6\n"                             # stops at the bad field
    "123456789.123456789,-0,.5\n"
    "9,8,7"                               # no trailing newline: still a line
)


@pytest.mark.skipif(not os.path.exists(LOADCHECK), reason="bin/loadcheck not built")
@pytest.mark.parametrize("case", ["tricky", "tricky_crlf", "cow_tr2", "cow_ref"])
def test_loader_and_writer_match_oracle(tmp_path, cloud_dir, case):
    if case.startswith("tricky"):
        src = tmp_path / "in.txt"
        text = TRICKY if case == "tricky" else TRICKY.replace("\n", "\r\n")
        src.write_bytes(text.encode())
        src = str(src)
    else:
        src = cloud_text_path(case, cloud_dir)
    out = str(tmp_path / "out.txt")
    r = subprocess.run([LOADCHECK, src, out], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    assert "[load] opening" in r.stderr and "[output] output file" in r.stderr       # src/load.cc:9,96
    ref = O.load_matrix(src)
    ref_out = str(tmp_path / "ref_out.txt")
    O.write_matrix(ref_out, ref)
    assert open(out).read() == open(ref_out).read()
    got = np.loadtxt(out, delimiter=",", skiprows=1, ndmin=2)
    assert got.shape == ref.shape


@pytest.mark.skipif(not os.path.exists(LOADCHECK), reason="bin/loadcheck not built")
def test_loader_missing_file_exits_2(tmp_path):
    r = subprocess.run([LOADCHECK, str(tmp_path / "nope.txt"), str(tmp_path / "o.txt")], capture_output=True, text=True)
    assert r.returncode == 2 and "could not be opened" in r.stderr                    # src/load.cc:11-14
