"""CPU oracle for the optical-flow training hot path.

TEST INFRASTRUCTURE ONLY.  Nothing under ``optical_flow_b200/`` may import this
package; only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s
``cpu_baseline`` / ``--impl reference`` legs use it, and there only as the
checker / reported CPU baseline, never as the thing shipped.

PARITY UNPINNED: the reference (xianlopez/optical_flow) is TensorFlow-2/Keras
code, TensorFlow is not installable in this environment (no network), and the
reference ships no tests, golden vectors or fixtures for this path
(SURVEY.md section 4, section 8c).  The oracle is therefore a line-by-line
restatement of the cited reference lines plus the TF defaults listed in
SURVEY.md F4-F7, cross-checked by two independent implementations
(``ref_np``: literal NumPy forward + hand-derived backward; ``ref_torch``:
torch-CPU autograd) and by the authored known-answer tests in
``tests/test_oracle_*.py``.
"""
