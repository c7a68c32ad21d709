"""CPU oracle for the models.py predict path -- TEST INFRASTRUCTURE ONLY.

This package restates, on the CPU, what the reference's hot path computes
(/root/reference/models.py:23-71): Pillow decode + NEAREST resize, Keras
``preprocess_input`` (caffe / tf mode), the Keras ResNet50 / InceptionV3
graphs, softmax and ``decode_predictions(top=5)``.

PARITY UNPINNED: the arithmetic of the reference lives in TensorFlow/Keras
(un-vendored, un-pinned, ~2.11 by date; ``pip install tensorflow`` at
/root/reference/test.py:61-63), which is not installed here and cannot be
installed offline, and the reference has no tests or numeric golden vectors
for this path.  The restatement follows the published
Keras application definitions; the only reference artefacts it is checked
against are the JSON shape and the (wnid, name) pairs of
/root/reference/download/output_*.json.  The front end (load_img) IS pinned:
it is checked bit-for-bit against Pillow itself, the same library the
reference calls.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
--impl reference legs may import this package.  The product package
(distributed-machine-learning_b200/) never does.
"""
