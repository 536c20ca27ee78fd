"""CPU oracle for the optical-flow training hot path.

TEST INFRASTRUCTURE ONLY.  Nothing under ``optical_flow_b200/`` may import this
package; only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s
``cpu_baseline`` / ``--impl reference`` legs use it, and there only as the
checker / reported CPU baseline, never as the thing shipped.

PARITY PINNED TO THE REFERENCE'S OWN PYTHON, NOT TO TENSORFLOW'S BINARIES: the reference
(xianlopez/optical_flow) is TensorFlow-2/Keras code, TensorFlow is not installable in this environment (no
network) and the reference ships no tests, golden vectors or fixtures for this path (SURVEY.md section 4,
section 8c).  What pins the oracle is tests/golden/: the reference's unmodified model.py, loss.py,
transformations.py, resnet/models.py, old/image_warp.py and old/loss.py executed over a torch-backed stand-in for
the ~40 TensorFlow calls they make (tests/golden/tf_shim, written from TF's documented op semantics and the
defaults listed in SURVEY.md F4-F7), with the generating scripts committed (make_golden.py, make_golden_geom.py;
make_golden_data.py uses the real cv2).  tests/test_a2_golden.py, tests/test_d2_geometry.py and
tests/test_d1_data_path.py hold every oracle function to those vectors.  Besides that the oracle is two
independent implementations (``ref_np*``: literal NumPy forward + hand-derived backward; ``ref_torch``:
torch-CPU autograd) that must agree, plus the authored known-answer tests of ``tests/test_a0_oracle.py``.
"""
