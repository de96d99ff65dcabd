"""CPU: decode gate (host logic) against the literal walk of read_image.py:57-128 in oracle/gate_ref.py."""
import random

from oracle import gate_ref
from video_edge_ai_proxy_b200.gate import (ACTIVITY_WINDOW_MS, REDIS_KEY_FRAME_ONLY_PREFIX, REDIS_LAST_ACCESS_PREFIX,
                                           DecodeGate, camera_active, poll_control)


def _product_walk(events):
    g = DecodeGate()
    groups, out = [], []
    for n, (kf, corrupt, kfo, q, nframes) in enumerate(events):
        if kf:
            groups = []
        groups.append((n, nframes))
        for o in g.on_packet(kf, corrupt, kfo, q):
            src, nf = groups[o.group_index]
            for f in range(nf):
                out.append(dict(source_packet=src, group_index=o.group_index, frame=f, packet=o.packet,
                                keyframe=o.keyframe, is_keyframe=o.is_keyframe, is_corrupt=o.is_corrupt))
                if o.first_frame_only:
                    break
                g.frame_published(kfo, q)
            g.packet_done()
    return out


def _script(rnd, n, gop, p_query, p_kfo):
    ev, q = [], None
    for i in range(n):
        if rnd.random() < p_query:
            q = (q or 1000) + rnd.randint(1, 50)
        ev.append((i % gop == 0, rnd.random() < .05, rnd.random() < p_kfo, q, rnd.choice([1, 1, 1, 2, 0])))
    return ev


def test_gate_matches_reference_walk_on_random_scripts():
    rnd = random.Random(3)
    for _ in range(300):
        ev = _script(rnd, rnd.randint(1, 120), rnd.choice([1, 5, 12, 30]), rnd.choice([0, .1, .5, 1]), rnd.choice([0, 0, .3, 1]))
        assert _product_walk(ev) == gate_ref.run(ev)


def test_keyframe_only_decodes_one_frame_per_gop():
    ev = [(i % 30 == 0, False, True, 5000 + i, 1) for i in range(90)]
    out = _product_walk(ev)
    assert [o["source_packet"] for o in out] == [0, 30, 60]
    assert [o["keyframe"] for o in out] == [1, 2, 3]


def test_no_query_decodes_keyframes_only_and_burst_catches_up():
    ev = [(i % 10 == 0, False, False, None, 1) for i in range(10)]
    assert [o["source_packet"] for o in _product_walk(ev)] == [0]
    # a query arriving at packet 6 forces packets 1..6 to be decoded in order (read_image.py:81-85)
    ev = [(i % 10 == 0, False, False, 777 if i >= 6 else None, 1) for i in range(8)]
    out = _product_walk(ev)
    assert [o["source_packet"] for o in out] == [0, 1, 2, 3, 4, 5, 6]
    assert [o["packet"] for o in out] == [0, 1, 2, 3, 4, 5, 6]
    # the same query is not served twice: packet 7 is skipped
    assert out == gate_ref.run(ev)


def test_activity_window():
    assert camera_active(1000, 1000 + ACTIVITY_WINDOW_MS - 1)
    assert not camera_active(1000, 1000 + ACTIVITY_WINDOW_MS)
    assert not camera_active(None, 5)


class _FakePipe:
    def __init__(self, store):
        self.store, self.ops = store, []

    def hgetall(self, k):
        self.ops.append(("h", k))

    def get(self, k):
        self.ops.append(("g", k))

    def execute(self):
        return [self.store.get(k, {} if t == "h" else None) for t, k in self.ops]


class _FakeRedis:
    def __init__(self, store):
        self.store, self.pipes = store, 0

    def pipeline(self):
        self.pipes += 1
        return _FakePipe(self.store)


def test_poll_control_reads_reference_keys_in_one_round_trip():
    r = _FakeRedis({REDIS_LAST_ACCESS_PREFIX + "cam1": {b"last_query": b"1604660000123", b"proxy_rtmp": b"1"},
                    REDIS_KEY_FRAME_ONLY_PREFIX + "cam1": b"True",
                    REDIS_KEY_FRAME_ONLY_PREFIX + "cam2": b"false"})
    ctl = poll_control(r, ["cam1", "cam2", "cam3"])
    assert r.pipes == 1
    assert ctl["cam1"].last_query == 1604660000123 and ctl["cam1"].key_frame_only and ctl["cam1"].proxy_rtmp
    assert ctl["cam2"].last_query is None and not ctl["cam2"].key_frame_only
    assert ctl["cam3"].last_query is None
