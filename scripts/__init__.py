"""Drop-in shim: `from scripts.als import ALS` and `from scripts.als_config import ...`
(the reference's import sites, tune_params.py:86-93, evaluate_models.py:95) resolve
to the B200 implementation.  If the reference checkout is importable too, its other
modules (create_folds, prepare_features, helpers, tune_params, evaluate_models) are
found through the extended package path and run unmodified against this ALS."""
import os as _os

_ref = _os.environ.get("B200ALS_REFERENCE_SCRIPTS", "/root/reference/scripts")
if _os.path.isdir(_ref) and _ref not in __path__:
    __path__.append(_ref)      # this repo's scripts/ stays first: als / als_config win
