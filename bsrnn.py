"""`from bsrnn import BSRNN` drop-in (reference train.py:11, infer.py:5, infer-streaming.py:6)."""
from speechseparation_b200.bsrnn import *  # noqa: F401,F403
from speechseparation_b200.bsrnn import BSRNN, generate_bandsplits  # noqa: F401
