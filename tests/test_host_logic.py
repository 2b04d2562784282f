"""Host-side logic that needs no GPU: the upload-group cuts of the host-buffer entry
(ptgs_b200.device.upload_group_cuts) and the launch-list summariser used for profiles/."""
import gzip
import os
import subprocess
import sys

import numpy as np

from ptgs_b200 import device as D

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _offsets(counts):
    return np.concatenate([[0], np.cumsum(counts)]).astype(np.int64)


def test_upload_group_cuts_equal_shares():
    rng = np.random.default_rng(3)
    counts = rng.integers(1000, 9000, size=37)
    off = _offsets(counts)
    for g in (1, 2, 4, 7, 37, 64):
        cuts = D.upload_group_cuts(off, 0, 37, g)
        assert cuts[0] == 0 and cuts[-1] == 37
        assert all(a <= b for a, b in zip(cuts, cuts[1:]))
        assert len(cuts) - 1 == min(g, 37)
        # a group never exceeds its share by more than one block
        sizes = [off[b] - off[a] for a, b in zip(cuts, cuts[1:])]
        assert sum(sizes) == off[-1]
        if g <= 7:
            assert max(sizes) <= off[-1] / g + counts.max()


def test_upload_group_cuts_shares_and_subrange():
    counts = np.full(20, 100)
    off = _offsets(counts)
    cuts = D.upload_group_cuts(off, 4, 14, [0.1, 0.4, 0.4, 0.1])     # 10 local blocks
    assert cuts == [4, 5, 9, 13, 14]
    assert D.upload_group_cuts(off, 4, 14, (1, 1)) == [4, 9, 14]       # shares are normalised
    # no local blocks: one empty group
    assert D.upload_group_cuts(off, 0, 0, 4) == [0, 0]
    # one huge block: the groups that fall inside it stay empty, nothing is lost
    off2 = _offsets([10, 10000, 10])
    cuts = D.upload_group_cuts(off2, 0, 3, 3)
    assert cuts[0] == 0 and cuts[-1] == 3 and all(a <= b for a, b in zip(cuts, cuts[1:]))


def test_launch_summary_only_filter(tmp_path):
    rows = ['"ID","Process ID","Process Name","Host Name","Kernel Name","Context","Stream","Block Size","Grid Size",'
            '"Device","CC","Section Name","Metric Name","Metric Unit","Metric Value"']
    def row(i, name, us):
        return '"%d","1","python","h","%s","1","7","(256, 1, 1)","(8, 1, 1)","0","10.0","Command line profiler metrics",' \
               '"gpu__time_duration.sum","us","%s"' % (i, name, us)
    rows += [row(0, "void ptgs::sweep_kernel<1>(ptgs::SweepArgs)", "3000.0"),
             row(1, "ptgs::cand_runs_kernel(const float *)", "800.0"),
             row(2, "void cub::DeviceRadixSortOnesweepKernel<x>(y)", "1200.0")]
    f = tmp_path / "l.csv.gz"
    with gzip.open(f, "wt") as fh:
        fh.write("==PROF== header line\n" + "\n".join(rows) + "\n")
    tool = os.path.join(ROOT, "tools", "launch_summary.py")
    full = subprocess.run([sys.executable, tool, str(f)], capture_output=True, text=True).stdout
    assert "launches 3" in full and "total kernel time 5.000 ms" in full
    only = subprocess.run([sys.executable, tool, str(f), "--only", "sweep_kernel|cand_"], capture_output=True,
                          text=True).stdout
    assert "launches 2" in only and "78.9%" in only and "DeviceRadixSort" not in only
