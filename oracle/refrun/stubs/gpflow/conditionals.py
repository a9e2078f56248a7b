def base_conditional(*args, **kwargs):
    raise NotImplementedError('imported by low_rank_calculations.py, never called on the exact path')
