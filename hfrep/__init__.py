"""Short import name for the package directory
`do-you-really-need-to-pay-2-20-hedge-fund-strategy-replication-via-machine-learning_b200`
(whose name is not a Python identifier).  `import hfrep` and `import hfrep.helper` return the
very same module objects."""
import importlib as _il
import sys as _sys

_LONG = "do-you-really-need-to-pay-2-20-hedge-fund-strategy-replication-via-machine-learning_b200"
_pkg = _il.import_module(_LONG)
for _name, _mod in list(_sys.modules.items()):
    if _name == _LONG or _name.startswith(_LONG + "."):
        _sys.modules["hfrep" + _name[len(_LONG):]] = _mod
_sys.modules["hfrep"] = _pkg
