"""Import alias: `from b200glue.pipeline import HostPipeline` (the package directory name has dashes)."""
import importlib

_p = importlib.import_module("cs566-project-lightglue_b200.pipeline")
HostPipeline = _p.HostPipeline
RESULT_KEYS = _p.RESULT_KEYS
pack_pairs = _p.pack_pairs
unpack_pred = _p.unpack_pred
group_by_length = _p.group_by_length
