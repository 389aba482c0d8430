"""Host logic of `BatchScheduler._admit` (chatterbox_b200/t3_batch.py) with a stub engine: queue order is kept, occupied slots
are skipped, and a run of consecutive free slots is filled by ONE `admit_many` call (so prompts of equal length share their
prefill pass) — the continuous-batching loop the reference's one-request-at-a-time server (src/server/server.py:189-299) lacks."""
from chatterbox_b200 import t3_batch as TB


class _Eng:
    n_slots = 6

    def __init__(self):
        self.calls = []

    def admit(self, slot, req):
        self.calls.append(("one", slot, [req]))

    def admit_many(self, slot0, reqs):
        self.calls.append(("many", slot0, list(reqs)))


def _sched(slots, queue):
    s = TB.BatchScheduler.__new__(TB.BatchScheduler)
    s.engine, s.slots, s.queue = _Eng(), list(slots), list(queue)
    return s


def test_runs_of_free_slots_are_admitted_together_in_queue_order():
    s = _sched([None, "x", None, None, "y", None], "abcde")
    s._admit()
    assert s.engine.calls == [("one", 0, ["a"]), ("many", 2, ["b", "c"]), ("one", 5, ["d"])]
    assert s.slots == ["a", "x", "b", "c", "y", "d"] and s.queue == ["e"]


def test_short_queue_and_full_engine():
    s = _sched([None] * 6, "ab")
    s._admit()
    assert s.engine.calls == [("many", 0, ["a", "b"])] and s.slots == ["a", "b", None, None, None, None] and s.queue == []
    s = _sched(["p"] * 6, "ab")
    s._admit()
    assert s.engine.calls == [] and s.queue == ["a", "b"]
    s = _sched([None] * 6, "")
    s._admit()
    assert s.engine.calls == []
