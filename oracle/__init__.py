"""CPU oracle for the ConceptAttention hot path.

TEST INFRASTRUCTURE ONLY.  This package is a plain fp32 CPU restatement of the
reference algorithm (manuragkhullar/ConceptAttention @ 401dc36).  Only
``tests/``, ``__graft_entry__.smoke()`` and the ``cpu_baseline`` / ``--impl
reference`` legs of ``bench.py`` may import it; ``conceptattention_b200`` (the
product) never does and fails loudly if its CUDA library is missing.

Parity status: PINNED.  ``tests/golden/make_golden.py`` imports the reference's
own Python modules from /root/reference (with stubs for the four third-party
packages absent from this image), runs them on seeded tiny-dimension inputs and
commits the outputs under ``tests/golden/*.npz``; ``tests/test_oracle_golden.py``
checks every oracle function against them.  Exception: ``sparsemax`` /
``entmax15`` come from the third-party ``entmax`` package, which is neither
vendored nor pinned by the reference and is not installed here -> those two
normalisers are "parity unpinned" (validated by their defining properties only).
"""
from .flux_math import *      # noqa: F401,F403
from .blocks import *         # noqa: F401,F403
from .heatmaps import *       # noqa: F401,F403
