"""Same-named module as the reference's `jax_llama/partition.py` (the code lives in `parallel/layout.py`)."""
from .parallel.layout import (Mesh, P, PartitionSpec, _get_partition_rules_llama, get_llama_param_partition_spec,  # noqa: F401
                              get_partition_spec, global_mesh_defined, with_named_sharding_constraint,
                              with_sharding_constraint)
