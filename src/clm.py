"""The reference's src/clm.py (WGAN-GP "criticizing language model" adversarial training) is disabled in every
shipped config (clm.enable: False) and is outside the accelerated path.  The module exists so
`from src.clm import CLM_wrapper` (solver.py:13) keeps importing; using it raises."""


class CLM_wrapper:
    def __init__(self, *args, **kwargs):
        raise NotImplementedError("CLM adversarial training is out of scope of this build; set clm.enable: False")
