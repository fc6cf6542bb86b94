"""Tensor-parallel layout: partition rules, mesh, shard / unshard of the canonical parameter tree.

The reference describes its layout as regex rules that map parameter paths to `PartitionSpec`s and
lets XLA/GSPMD do the rest (/root/reference/jax_llama/partition.py:16-81).  Here the same rule table
is the single source of truth, but it is *executed* explicitly: `shard_tree` cuts the canonical tree
(SURVEY §2.2 names, kernels in `[in, out]` layout) into the per-rank slices a Megatron-style TP engine
loads on one GPU each, and `unshard_tree` is its inverse (used by tests and checkpoint export).

Axis names follow the reference: ``"dp"`` (replica axis) and ``"mp"`` (tensor/model-parallel axis).
"""
from __future__ import annotations

import re
from dataclasses import dataclass, field
from typing import Any, Callable, Dict, Iterable, List, Mapping, Optional, Sequence, Tuple, Union

import numpy as np
import torch

Tree = Dict[str, Any]
Path = Tuple[str, ...]


class PartitionSpec(tuple):
    """One entry per tensor dimension: an axis name, a tuple of axis names, or None (replicated)."""

    def __new__(cls, *axes: Union[None, str, Tuple[str, ...]]):
        return super().__new__(cls, axes)

    def __repr__(self) -> str:
        return "PartitionSpec(" + ", ".join(repr(a) for a in self) + ")"

    def axis_of(self, name: str) -> Optional[int]:
        """Dimension index sharded over mesh axis `name`, or None."""
        for dim, a in enumerate(self):
            if a == name or (isinstance(a, tuple) and name in a):
                return dim
        return None


P = PartitionSpec


@dataclass(frozen=True)
class Mesh:
    """Logical device mesh: ordered axis names with sizes, e.g. Mesh({'dp': 1, 'mp': 8})."""

    shape: Mapping[str, int] = field(default_factory=lambda: {"dp": 1, "mp": 1})

    @property
    def axis_names(self) -> Tuple[str, ...]:
        return tuple(self.shape.keys())

    def size(self, axis: Optional[str] = None) -> int:
        if axis is None:
            return int(np.prod(list(self.shape.values())))
        return int(self.shape[axis])

    # `with mesh:` makes it the ambient mesh, the way `with jax.sharding.Mesh(...)` does for the reference's
    # `global_mesh_defined()` / `with_sharding_constraint` (partition.py:83-93)
    def __enter__(self) -> "Mesh":
        _MESH_STACK.append(self)
        return self

    def __exit__(self, *exc: Any) -> None:
        _MESH_STACK.pop()

    def coords(self, rank: int) -> Dict[str, int]:
        """Row-major rank → per-axis coordinate (last axis fastest, like jax's device mesh)."""
        out: Dict[str, int] = {}
        for name in reversed(self.axis_names):
            out[name] = rank % self.shape[name]
            rank //= self.shape[name]
        return out


_MESH_STACK: List["Mesh"] = []


def global_mesh_defined() -> bool:
    """True inside a `with Mesh(...):` block (partition.py:83-86; the reference asks jax for its ambient mesh)."""
    return bool(_MESH_STACK)


def current_mesh() -> Optional["Mesh"]:
    return _MESH_STACK[-1] if _MESH_STACK else None


# --------------------------------------------------------------------------- rule tables
Rule = Tuple[Tuple[str, ...], PartitionSpec]

_TP_RULES: List[Rule] = [
    (("transformer", "wte", "embedding"), P("mp", None)),          # vocab-parallel embedding
    (("attention", "(wq|wk|wv)", "kernel"), P(None, "mp")),        # column-parallel
    (("attention", "wo", "kernel"), P("mp", None)),                # row-parallel
    (("feed_forward", "(w1|w3)", "kernel"), P(None, "mp")),        # column-parallel
    (("feed_forward", "w2", "kernel"), P("mp", None)),             # row-parallel
    (("(attention_norm|ffn_norm)", "kernel"), P(None)),            # replicated
    (("transformer", "ln_f", "kernel"), P(None)),
    (("lm_head", "kernel"), P(None, "mp")),                        # vocab-parallel head
]

_FSDP_RULES: List[Rule] = [
    (("transformer", "wte", "embedding"), P("mp", "dp")),
    (("attention", "(wq|wk|wv)", "kernel"), P("dp", "mp")),
    (("attention", "wo", "kernel"), P("mp", "dp")),
    (("feed_forward", "(w1|w3)", "kernel"), P("dp", "mp")),
    (("feed_forward", "w2", "kernel"), P("mp", "dp")),
    (("(attention_norm|ffn_norm)", "kernel"), P(None)),
    (("transformer", "ln_f", "kernel"), P(None)),
    (("lm_head", "kernel"), P("dp", "mp")),
]


def partition_rules(fsdp: bool = False) -> List[Rule]:
    """TP table (the one exercised) or the 2-D FSDP table (API parity; partition.py:44-61)."""
    return list(_FSDP_RULES if fsdp else _TP_RULES)


# --------------------------------------------------------------------------- tree helpers
_get_partition_rules_llama = partition_rules        # the reference's (private) name, partition.py:43


def flatten_tree(tree: Mapping[str, Any], prefix: Path = ()) -> Dict[Path, Any]:
    flat: Dict[Path, Any] = {}
    for k, v in tree.items():
        if isinstance(v, Mapping):
            flat.update(flatten_tree(v, prefix + (str(k),)))
        else:
            flat[prefix + (str(k),)] = v
    return flat


def unflatten_tree(flat: Mapping[Path, Any]) -> Tree:
    tree: Tree = {}
    for path, v in flat.items():
        node = tree
        for k in path[:-1]:
            node = node.setdefault(k, {})
        node[path[-1]] = v
    return tree


def tree_map(fn: Callable[..., Any], tree: Mapping[str, Any], *rest: Mapping[str, Any]) -> Tree:
    flat = flatten_tree(tree)
    others = [flatten_tree(r) for r in rest]
    return unflatten_tree({p: fn(v, *(o[p] for o in others)) for p, v in flat.items()})


def _window_match(patterns: Sequence[str], path: Path) -> bool:
    """True if the regex sequence fully matches some contiguous window of path components."""
    n = len(patterns)
    compiled = [re.compile(p) for p in patterns]
    for start in range(len(path) - n + 1):
        if all(c.fullmatch(path[start + j]) for j, c in enumerate(compiled)):
            return True
    return False


def get_partition_spec(tree: Mapping[str, Any], rules: Iterable[Rule]) -> Tree:
    """First matching rule wins; a leaf no rule matches is an error (as in the reference)."""
    rules = list(rules)
    out: Dict[Path, PartitionSpec] = {}
    for path in flatten_tree(tree):
        for patterns, spec in rules:
            if _window_match(patterns, path):
                out[path] = spec
                break
        else:
            raise ValueError(f"Incomplete partition spec: no rule matches {'/'.join(path)}")
    return unflatten_tree(out)


def get_llama_param_partition_spec(params: Mapping[str, Any], fsdp: bool = False) -> Tree:
    return get_partition_spec(params, partition_rules(fsdp))


# --------------------------------------------------------------------------- executing the layout
def _slice_dim(x: Any, dim: int, index: int, parts: int, what: str) -> Any:
    n = x.shape[dim]
    if n % parts:
        raise ValueError(f"{what}: dimension {dim} of size {n} is not divisible by {parts}")
    step = n // parts
    sl = [slice(None)] * x.ndim
    sl[dim] = slice(index * step, (index + 1) * step)
    return x[tuple(sl)]


def shard_leaf(x: Any, spec: PartitionSpec, mesh: Mesh, rank: int, what: str = "param") -> Any:
    coords = mesh.coords(rank)
    for dim, axes in enumerate(spec):
        if axes is None:
            continue
        for axis in ((axes,) if isinstance(axes, str) else axes):
            if mesh.size(axis) > 1:
                x = _slice_dim(x, dim, coords[axis], mesh.size(axis), what)
    return x


def shard_tree(tree: Mapping[str, Any], mesh: Mesh, rank: int, fsdp: bool = False) -> Tree:
    """Slice every leaf of the canonical tree for `rank` of `mesh` according to the rule table."""
    specs = flatten_tree(get_llama_param_partition_spec(tree, fsdp))
    flat = flatten_tree(tree)
    return unflatten_tree({p: shard_leaf(v, specs[p], mesh, rank, "/".join(p)) for p, v in flat.items()})


def unshard_tree(shards: Sequence[Mapping[str, Any]], mesh: Mesh, fsdp: bool = False) -> Tree:
    """Inverse of `shard_tree` over all ranks of `mesh` (row-major rank order)."""
    if len(shards) != mesh.size():
        raise ValueError(f"expected {mesh.size()} shards, got {len(shards)}")
    specs = flatten_tree(get_llama_param_partition_spec(shards[0], fsdp))
    flats = [flatten_tree(s) for s in shards]
    out: Dict[Path, Any] = {}
    for path, spec in specs.items():
        pieces = {tuple(mesh.coords(r).items()): flats[r][path] for r in range(mesh.size())}

        def assemble(fixed: Dict[str, int], axes_left: List[Tuple[int, str]]) -> Any:
            if not axes_left:
                coords = {a: fixed.get(a, 0) for a in mesh.axis_names}
                rank_key = tuple((a, coords[a]) for a in reversed(mesh.axis_names))
                return pieces[rank_key]
            dim, axis = axes_left[0]
            parts = [assemble({**fixed, axis: i}, axes_left[1:]) for i in range(mesh.size(axis))]
            if isinstance(parts[0], torch.Tensor):
                return torch.cat(parts, dim=dim)
            return np.concatenate(parts, axis=dim)

        axes: List[Tuple[int, str]] = []
        for dim, a in enumerate(spec):
            if a is None:
                continue
            for axis in ((a,) if isinstance(a, str) else a):
                if mesh.size(axis) > 1:
                    axes.append((dim, axis))
        out[path] = assemble({}, axes)
    return unflatten_tree(out)


def check_tp_divisibility(config: Any, tp: int) -> None:
    """Raise early, with a readable message, if a model cannot be cut `tp` ways."""
    problems = []
    if config.num_attention_heads % tp:
        problems.append(f"num_attention_heads={config.num_attention_heads}")
    if config.num_key_value_heads % tp:
        problems.append(f"num_key_value_heads={config.num_key_value_heads}")
    if config.intermediate_size % tp:
        problems.append(f"intermediate_size={config.intermediate_size}")
    if config.vocab_size % tp:
        problems.append(f"vocab_size={config.vocab_size}")
    if problems:
        raise ValueError(f"tensor-parallel degree {tp} does not divide: " + ", ".join(problems))


# --------------------------------------------------------------------------- constraint wrappers
def with_sharding_constraint(x: Any, axis_resources: Any) -> Any:
    """API parity with partition.py:88-93.

    In the reference this is a hint to GSPMD and inert outside a mesh context.  Our runtime has no
    tracing compiler: activations are laid out explicitly by the engine, so the call returns `x`
    unchanged after validating that the requested spec is applicable to its rank.
    """
    spec = axis_resources[1] if isinstance(axis_resources, tuple) and len(axis_resources) == 2 \
        and isinstance(axis_resources[0], Mesh) else axis_resources
    if isinstance(spec, PartitionSpec) and hasattr(x, "ndim") and len(spec) > x.ndim:
        raise ValueError(f"{spec} has more entries than the array has dimensions ({x.ndim})")
    return x


def with_named_sharding_constraint(x: Any, mesh: Optional[Mesh], partition_spec: PartitionSpec) -> Any:
    """API parity with partition.py:95-98: no-op without a mesh; validates batch divisibility."""
    if mesh is None:
        return x
    for dim, a in enumerate(partition_spec):
        if a is None or not hasattr(x, "shape") or dim >= len(x.shape):
            continue
        for axis in ((a,) if isinstance(a, str) else a):
            if x.shape[dim] % mesh.size(axis):
                raise ValueError(f"dimension {dim} of size {x.shape[dim]} not divisible by mesh axis "
                                 f"{axis!r} of size {mesh.size(axis)}")
    return with_sharding_constraint(x, (mesh, partition_spec))
