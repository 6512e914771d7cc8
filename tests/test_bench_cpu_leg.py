"""The CPU legs of bench.py (`cpu_baseline`, `--impl reference`): the oracle
port run as independent volume replicas on every host core must do exactly the
work of the two-thread single replica (same keyframes, same number of
`_update_voxel` events) -- only the wall time may differ."""
import bench


def test_replicas_do_the_same_work():
    from oracle import oracle as orc

    orc.build()
    kf = bench.gen_batch(bench.TARGET, 0, 2)
    ev1, _ = bench.cpu_fuse(kf, range(2), threads=1)   # sequential, one replica
    ev2, _ = bench.cpu_fuse(kf, range(2), threads=2)   # one replica, two threads
    ev4, _ = bench.cpu_fuse(kf, range(2), threads=4)   # two replicas, one keyframe each
    assert ev1 == ev2 == ev4 > 1_000_000
    assert bench.host_threads() >= 1


def test_both_arms_share_one_config_dict():
    """`config` is what the driver compares between the GPU arm and the
    reference arm (`same_config`): it must not depend on N or on the arm."""
    import argparse

    a = argparse.Namespace(kf_per_launch=40, launches_per_step=15, passes_per_step=4, gpus=1)
    b = argparse.Namespace(kf_per_launch=40, launches_per_step=15, passes_per_step=4, gpus=8)
    assert bench.workload_config(a) == bench.workload_config(b)
    assert "1024^3 @ 5 mm" in bench.workload_config(a)["workload"]


def test_packed_layout_keeps_tma_alignment():
    from mast3r_slam_quality_dualtsdf_b200.tsdf.global_volume import packed_layout

    for share in (1, 5, 10, 20, 40):
        lay = packed_layout(share, bench.H, bench.W, bench.GH, bench.GW)
        assert lay["seg_len"] % 4 == 0 and lay["C"] % 4 == 0 and lay["pose"] % 4 == 0
        assert lay["pose"] + share * 8 <= lay["seg_len"]
        assert lay["q"] + share * bench.GH * bench.GW <= lay["pose"]
