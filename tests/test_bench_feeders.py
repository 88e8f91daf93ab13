"""bench.py's frame feeders (CPU): the protocol by which frames reach an engine with encoder look-ahead -- frame f as a
view, the stack of the next B frames handed over with the call that consumes the last frame encoded ahead, at most
three upload buffers alive (FrameStreamer's depth) -- simulated with a mock streamer and a mock engine."""
import collections
import importlib.util
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def bench():
    argv = sys.argv
    sys.argv = ["bench.py"]
    try:
        spec = importlib.util.spec_from_file_location("bench_under_test", os.path.join(ROOT, "bench.py"))
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
    finally:
        sys.argv = argv
    return mod


class _Stack(list):
    @property
    def shape(self):
        return (len(self),)


class _Streamer:
    depth = 3

    def __init__(self):
        self.staged = collections.deque()
        self.alive = self.max_alive = 0

    def stage_stack(self, imgs):
        self.staged.append(list(imgs))
        self.alive += 1
        self.max_alive = max(self.max_alive, self.alive)

    def acquire(self):
        return _Stack(self.staged.popleft())

    def release(self, t):
        self.alive -= 1


@pytest.mark.parametrize("la", [1, 2, 3, 4, 6])
@pytest.mark.parametrize("n_frames", [40, 61, 97])
def test_host_feeder_protocol(bench, la, n_frames):
    limit = 2 + la * (-(-(n_frames - 1) // la))
    src = list(range(limit + 8))
    fs = _Streamer()
    feed = bench.HostFeeder(fs, src, la, limit)
    ahead, encoded = 0, []
    for f in range(1, n_frames):
        assert feed.frame(f) == [f]
        if ahead == 0:
            ahead = 1                                  # nothing encoded ahead: the engine encodes `img` itself
            encoded.append(f)
        nxt = None
        if ahead <= 1 and f + 1 < limit:
            nxt = feed.stack_after(f)
            assert list(nxt) == list(range(f + 1, f + 1 + la))     # always a full stack
            encoded += list(nxt)
        ahead -= 1
        if nxt is not None:
            ahead += len(nxt)
    assert fs.max_alive <= fs.depth
    assert encoded[:n_frames - 1] == list(range(1, n_frames))      # every frame encoded once, in order


def test_lookahead_divides_steps(bench):
    """The timed region holds exactly `steps` frames of encoder work when the look-ahead divides `steps`: stacks are
    launched at calls f = 1 (mod la), so any window of `steps` consecutive calls contains steps / la launches."""
    for la in (1, 2, 3, 4, 6):
        for first in (22, 23, 24, 31):
            steps = 60
            launches = sum(1 for f in range(first, first + steps) if f % la == 1 % la)
            assert launches * la == steps
