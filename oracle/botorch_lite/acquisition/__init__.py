from . import analytic
