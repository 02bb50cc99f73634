import time

from fedrec_b200.utils.health import FaultInjector, Heartbeat


class _Store(dict):
    def set(self, k, v):
        self[k] = v.encode()

    def get(self, k):
        return self[k]

    def check(self, ks):
        return all(k in self for k in ks)


def test_heartbeat_flags_stalled_rank():
    store = _Store()
    a = Heartbeat(store, 0, 3, stale_s=5.0)
    b = Heartbeat(store, 1, 3)
    a.beat(); b.beat()
    t0 = time.time()
    assert a.check_once(t0) == []
    b.beat()
    assert a.check_once(t0 + 4) == []                  # rank 1 moved; rank 2 silent but inside the window
    b.beat()
    assert a.check_once(t0 + 10) == [2]                # rank 2 never reported
    assert a.check_once(t0 + 30) == [1, 2]             # rank 1 stopped too


def test_fault_injector_delay():
    f = FaultInjector("delay:1:3:0.05")
    t = time.time()
    f.maybe_fire(0, 3); f.maybe_fire(1, 2)
    assert time.time() - t < 0.04
    f.maybe_fire(1, 3)
    assert time.time() - t >= 0.05
