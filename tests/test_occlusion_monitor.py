"""CPU: the occlusion hysteresis state machine (fused_depth_map.py:2325-2337, 2515-2571) replayed against a
line-by-line transcription of the reference's loop body on recorded detector sequences."""
import numpy as np

from stereovision_b200.fusion import OcclusionMonitor


def reference_loop(detections, use_stereo=True, hysteresis=5, interval=2):
    """fused_depth_map.py:2515-2571 with the prints removed; detections[k] is what detect_camera_occlusion
    returns on frame k.  Yields USE_STEREO after each frame."""
    state, counter = "none", 0
    for frame_counter, new_state in enumerate(detections):
        if frame_counter % interval == 0:
            if new_state == state:
                counter = 0
            else:
                counter += 1
                if counter >= hysteresis:
                    state = new_state
                    counter = 0
                    if state in ["left", "right"]:
                        if use_stereo:
                            use_stereo = False
                    elif state in ["none", "both"]:
                        if not use_stereo:
                            use_stereo = True
        yield use_stereo, state


def test_left_occluded_sequence_switches_stereo_off_after_five_checks():
    seq = ["none"] * 4 + ["left"] * 20 + ["none"] * 20
    mon = OcclusionMonitor(detect=lambda l, r, occlusion_threshold: (seq[mon_frame[0]], 0.9, 0.1))
    mon_frame = [0]
    got = []
    for k in range(len(seq)):
        mon_frame[0] = k
        got.append((mon.update(k, None, None), mon.state))
    assert got == list(reference_loop(seq))
    # frames 4, 6, 8, 10, 12 are the five consecutive differing checks: stereo goes off on frame 12
    assert [g[0] for g in got[:12]] == [True] * 12 and got[12] == (False, "left")
    # recovery: 'none' from frame 24 on, checks at 24 .. 32 -> back on at 32
    assert got[31] == (False, "left") and got[32] == (True, "none")
    assert mon.left_score == 0.9 and mon.right_score == 0.1


def test_random_sequences_match_the_reference_loop():
    rng = np.random.default_rng(4)
    states = ["none", "left", "right", "both"]
    for it in range(50):
        runs = []
        while len(runs) < 120:
            runs += [states[int(rng.integers(0, 4))]] * int(rng.integers(1, 16))
        mon = OcclusionMonitor(use_stereo=bool(it % 2))
        got = []
        for k, s in enumerate(runs):
            if k % mon.check_interval == 0:
                mon.observe(s)
            got.append((mon.use_stereo, mon.state))
        assert got == list(reference_loop(runs, use_stereo=bool(it % 2)))


def test_both_cameras_covered_keeps_stereo_on():
    mon = OcclusionMonitor()
    for _ in range(5):
        mon.observe("both")
    assert mon.state == "both" and mon.use_stereo
    for _ in range(4):
        mon.observe("right")
    mon.observe("both")                       # an agreeing check resets the count
    for _ in range(4):
        mon.observe("right")
    assert mon.use_stereo and mon.state == "both"
    mon.observe("right")
    assert not mon.use_stereo and mon.state == "right"
