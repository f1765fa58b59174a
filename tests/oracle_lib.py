"""The ctypes loader of the CPU oracle lives with the oracle (oracle/oracle_lib.py); re-exported for the tests."""
from oracle.oracle_lib import *  # noqa: F401,F403
from oracle.oracle_lib import ORACLE_DIR, Oracle, load  # noqa: F401
