"""Import-only stand-in for `ray`, used by oracle/make_golden.py so that the reference script
/root/reference/ray-jobs/pytorch_llm_ray.py (which does `import ray` at module top, ref :12-18) can be imported
in this container and its model/dataset classes exercised on stock torch. No Ray behaviour is emulated."""
from . import train, cloudpickle  # noqa: F401
