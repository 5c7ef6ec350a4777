"""Measurement switch: with the stub on, the bulk data-plane transfers of a training step (pipeline activation /
gradient transfers, the data-parallel gradient reduction and parameter all-gather) are skipped. The step then computes
garbage at exactly the same compute cost, and `step time - stubbed step time` is the communication (and the waiting it
causes) that the real step does not hide: the "exposed comm / step" half of the headline metric.
Only the benchmark turns it on."""
_on = False


class _Done:
    """Stand-in for a finished torch.distributed Work."""

    def wait(self):
        return True

    def is_completed(self):
        return True


DONE = _Done()


def enabled():
    return _on


def set_enabled(flag):
    global _on
    prev, _on = _on, bool(flag)
    return prev
