"""``from model import Model`` keeps working inside this variant directory."""
import os, sys
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))
from ddl_b200.models.compat import Model  # noqa: F401,E402
