"""Drop-in `utils` package for the unchanged Fairify drivers.

The drivers do `from utils.input_partition import *`, `from utils.verif_utils import *`,
`from utils.prune import *` and `import_module('utils.<MODEL>-Model-Functions')`
(reference src/AC/Verify-AC.py:19-22, 86-90) with the repository root on sys.path.  Putting THIS repository
first on sys.path routes the pre-solver stage (partition grid, sampling, activation profile, interval
bounds, masks, logits) through libfairify_b200.so.  Everything that is not on that path -- the Z3 encoders and
property builders, the dataset loaders -- is re-exported from a reference checkout when one is available
(FAIRIFY_REFERENCE_ROOT), because the north star keeps that code on the host, unchanged.

The 53 `utils/<MODEL>-Model-Functions.py` files of the reference all define the same generic `layer_net` /
`net` / `z3_net`; here one module (utils/_model_functions.py) serves every such import through a finder.
"""
import importlib.abc
import importlib.util
import os
import sys


class _ModelFunctionsFinder(importlib.abc.MetaPathFinder):
    def find_spec(self, fullname, path, target=None):
        if fullname.startswith("utils.") and fullname.endswith("-Model-Functions"):
            src = os.path.join(os.path.dirname(__file__), "_model_functions.py")
            return importlib.util.spec_from_file_location(fullname, src)
        return None


if not any(isinstance(f, _ModelFunctionsFinder) for f in sys.meta_path):
    sys.meta_path.insert(0, _ModelFunctionsFinder())
