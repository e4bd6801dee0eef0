"""bench.py's command-line contract on the CPU: both arms print the same `config` for the same arguments (the driver
compares them), defaults finish quickly, the step holds two concurrent launch groups, the host entry point's group size."""
import sys

import bench
from applied_image_processing_b200 import localized_style_transfer as lst


def _args(*argv):
    old = sys.argv
    sys.argv = ["bench.py", *argv]
    try:
        return bench.parse_args()
    finally:
        sys.argv = old


def test_both_arms_share_the_config_dict():
    for argv in ([], ["--gpus", "8"], ["--workload", "1080p", "--frames", "128", "--gpus", "8"], ["--workload", "16k", "--gpus", "4"],
                 ["--workload", "16k", "--bins", "1024"], ["--workload", "labels"], ["--kind", "N", "--sty-short", "512"]):
        ours, ref = _args(*argv), _args(*argv, "--impl", "reference")
        assert bench.config_dict(ours, ours.gpus) == bench.config_dict(ref, ref.gpus), argv
        cfg = bench.config_dict(ours, ours.gpus)
        assert "workload" in cfg and "model" not in cfg and "l2" in cfg


def test_defaults():
    a = _args()
    assert (a.gpus, a.workload, a.frames, a.slots, a.lanes) == (1, "4k", 16, 8, 2) and a.steps >= 3 and a.warmup >= 3
    assert "BASELINE.json configs[1]" in bench.workload_name(a)
    assert _args("--workload", "16k").lanes == 1 and _args("--bins", "1024").lanes == 1
    assert "BINNED APPROXIMATION" in bench.workload_name(_args("--bins", "1024"))


def test_host_group_size_and_lanes():
    assert [lst.host_group_size(n) for n in (1, 8, 16, 32, 1024)] == [1, 1, 1, 1, 1]                     # 4K frames: one per group
    assert [lst.host_group_size(n, 14_515_200) for n in (1, 3, 8, 16, 32, 128)] == [1, 1, 2, 4, 4, 4]     # 1080p frames
    assert lst.LANES == 2
