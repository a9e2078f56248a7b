"""``from tensorflow.contrib import stateless`` (low_rank_calculations.py:6): imported by the reference, never called on the exact path."""
stateless = None
