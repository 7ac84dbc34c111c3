"""CPU oracle for the batched image `predict` path — TEST INFRASTRUCTURE ONLY.

Nothing under oracle/ is part of the product: only tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs may import it, and only as the checker or the timed CPU
baseline.  The product path (mlmicroservicetemplate_b200/) never imports this package and has no CPU
fallback.

PARITY UNPINNED BY THE REFERENCE.  CodyRichter/MLMicroserviceTemplate ships no tests, golden vectors
or known-answer fixtures, and its predict() (src/model/model.py:16-31) performs no arithmetic: it
opens the upload with PIL (model.py:23) and returns a constant dict (model.py:25-31).  The arithmetic
the north star adds lives in third-party libraries that the reference does not even pin
(requirements.txt:1-11 lists `pillow` unpinned and no torch/torchvision at all).  The oracle is
therefore the reference's predict() *contract* filled in with those libraries at the versions
installed in this image:

    torch 2.11.0 (CPU: oneDNN/MKL), torchvision 0.26.0, Pillow 12.2.0

and it is pinned by (1) the tolerances stated in BASELINE.json's north_star, (2) fixed seeds, and
(3) the golden fixtures under tests/golden/ that were generated in this container by importing
those libraries and the reference server itself (tests/golden/make_golden.py).

Modules:
  model.py         init()/predict() with the reference's signatures (src/model/model.py:6,16), CPU fp32.
  preprocess_np.py numpy restatement of the antialiased resize + normalise (validated bit-exact
                   against torch's F.interpolate on this build) — the spec the CUDA kernel follows.
  images.py        deterministic structured synthetic images.
"""
