"""A/B on one box: bench.py with ('with') or without ('without') a barrier before the slot is overwritten.
usage (under torchrun): ab_slot_barrier.py with|without <bench.py arguments>"""
import os
import runpy
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from milnce_b200 import utils  # noqa: E402

if sys.argv[1] == "with":
    _orig = utils.PeerRing.next_slot

    def _next_slot(self):
        self.hdl.barrier(channel=0)
        return _orig(self)
    utils.PeerRing.next_slot = _next_slot
sys.argv = [os.path.join(ROOT, "bench.py")] + sys.argv[2:]
runpy.run_path(sys.argv[0], run_name="__main__")
