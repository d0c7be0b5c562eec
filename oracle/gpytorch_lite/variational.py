"""Import-only stubs: fs_mol/utils/gp_utils.py:4 imports these for the ADKT-VGP ablation (OUT OF SCOPE)."""


class CholeskyVariationalDistribution:
    def __init__(self, *a, **k):
        raise NotImplementedError("variational GP is out of scope (SURVEY.md section 2 row 2)")


class UnwhitenedVariationalStrategy:
    def __init__(self, *a, **k):
        raise NotImplementedError("variational GP is out of scope (SURVEY.md section 2 row 2)")
