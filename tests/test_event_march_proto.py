"""The event-driven form of tracing algorithm 3 that DESIGN.md plans for large scenes (tools/event_march_proto.py):
deferring the per-object subtractions and replaying them only for the candidates of the minimum must reproduce the
eager form bit for bit (final length, trip count, number of SDF evaluations).  CPU only, numpy."""
import importlib.util
import os

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_event_driven_march_equals_eager_march_bit_for_bit():
    spec = importlib.util.spec_from_file_location("event_march_proto", os.path.join(ROOT, "tools", "event_march_proto.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    bad, tot, n = mod.run(rays=3, seed=11)
    assert bad == 0
    assert sum(t["rays"] for t in tot.values()) == 3 * n
    # the point of the design: almost no deferred work is ever replayed
    trips = sum(t["trips"] for t in tot.values())
    assert sum(t["replayed_objects"] for t in tot.values()) < 3 * trips


def test_calendar_queue_march_equals_eager_march_bit_for_bit():
    """Second step of the same design (tools/event_march_buckets.py): the objects live in a calendar queue over the absolute
    ray length and a trip examines only the buckets within reach."""
    import sys
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    spec = importlib.util.spec_from_file_location("event_march_buckets", os.path.join(ROOT, "tools", "event_march_buckets.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    bad, tot, n = mod.run(rays=2, delta=0.5, B=4096, seed=7)
    assert bad == 0
    trips = sum(t["trips"] for t in tot.values())
    examined = sum(t["examined"] for t in tot.values())
    assert examined < 400 * trips          # against 10,000 objects touched per trip by the eager form
