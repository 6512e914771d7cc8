"""The CPU legs of bench.py (`cpu_baseline`, `--impl reference`): the oracle
port run as independent volume replicas on every host core must do exactly the
work of the two-thread single replica (same keyframes, same number of
`_update_voxel` events) -- only the wall time may differ."""
import bench


def test_replicas_do_the_same_work():
    from oracle import oracle as orc

    orc.build()
    scene, dim, orbit_r = bench.scene_for(1)
    kf = bench.gen_batch(scene, orbit_r, 0, 2)
    ev1, _ = bench.cpu_fuse(kf, range(2), threads=1)   # sequential, one replica
    ev2, _ = bench.cpu_fuse(kf, range(2), threads=2)   # one replica, two threads
    ev4, _ = bench.cpu_fuse(kf, range(2), threads=4)   # two replicas, one keyframe each
    assert ev1 == ev2 == ev4 > 1_000_000
    assert bench.host_threads() >= 1
