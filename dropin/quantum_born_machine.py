"""Import shim: `import quantum_born_machine` resolves to tensornetworks_b200.quantum_born_machine when this directory is on
sys.path, so scripts written against the reference's flat module layout run unchanged."""
import os as _os
import sys as _sys

_sys.path.insert(0, _os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))))
from tensornetworks_b200.quantum_born_machine import *  # noqa: E402,F401,F403
from tensornetworks_b200 import quantum_born_machine as _impl  # noqa: E402

globals().update({k: v for k, v in vars(_impl).items() if not k.startswith("__")})
