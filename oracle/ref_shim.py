"""TEST INFRASTRUCTURE ONLY -- never imported by the product path.

Loads the reference's own Ray Data exchange code (pure Python over pyarrow /
numpy) from ``/root/reference/python`` WITHOUT the compiled Ray core, so that

* ``tests/golden/make_golden.py`` can generate golden input/output vectors by
  running the reference itself, and
* the ``-m "not gpu"`` tests can check ``oracle/exchange_oracle.py`` (the
  restatement that travels to the GPU box) against the real reference whenever
  ``/root/reference`` is present.

``/root/reference`` does not exist on the GPU box: nothing that runs there
(``-m gpu`` tests, ``smoke()``, ``bench.py``) may call :func:`load`.

How it works: four module prefixes are missing in this
container -- ``ray._raylet`` (Cython extension), ``ray.core.generated`` and
``ray.core.src`` (bazel-built protobufs / binaries) and ``colorama`` (an
uninstalled dependency of ``ray._private.ray_logging``).  A ``sys.meta_path``
finder answers those with ``MagicMock``-backed package modules.  Two call sites
that would otherwise start a cluster are patched:
``ray.runtime_context.get_runtime_context`` (used by ``BlockExecStats``,
python/ray/data/block.py:156) and ``SizeEstimator._real_size``
(python/ray/data/_internal/size_estimator.py:55-63, which calls ``ray.put``).
"""
from __future__ import annotations

import importlib.abc
import importlib.machinery
import os
import sys
import types
from unittest import mock

REFERENCE_ROOT = os.environ.get("XCHG_REFERENCE_ROOT", "/root/reference")
_PY_ROOT = os.path.join(REFERENCE_ROOT, "python")
_MOCK_PREFIXES = ("ray._raylet", "ray.core.generated", "ray.core.src", "colorama")

_loaded = None


def available() -> bool:
    return os.path.isdir(os.path.join(_PY_ROOT, "ray", "data"))


class _MockModule(mock.MagicMock):
    """A MagicMock that can stand in for a package module."""

    __path__: list = []
    __all__: list = []


class _MockLoader(importlib.abc.Loader):
    def create_module(self, spec):
        m = _MockModule(name=spec.name)
        m.__name__ = spec.name
        m.__spec__ = spec
        m.__loader__ = self
        m.__path__ = []
        m.__all__ = []
        return m

    def exec_module(self, module):
        return None


class _MockFinder(importlib.abc.MetaPathFinder):
    def find_spec(self, fullname, path=None, target=None):
        for p in _MOCK_PREFIXES:
            if fullname == p or fullname.startswith(p + "."):
                return importlib.machinery.ModuleSpec(
                    fullname, _MockLoader(), is_package=True
                )
        return None


class _FakeRuntimeContext:
    def get_node_id(self):
        return "oracle-node"


def load() -> types.SimpleNamespace:
    """Import the reference hot-path modules; returns a namespace of them."""
    global _loaded
    if _loaded is not None:
        return _loaded
    if not available():
        raise RuntimeError(
            f"reference not present at {REFERENCE_ROOT}; the shim only works in "
            "the build container"
        )
    if _PY_ROOT not in sys.path:
        sys.path.insert(0, _PY_ROOT)
    if not any(isinstance(f, _MockFinder) for f in sys.meta_path):
        sys.meta_path.insert(0, _MockFinder())

    import ray  # noqa: F401  (reference's package, not an installed wheel)
    import ray.runtime_context as rc

    rc.get_runtime_context = lambda: _FakeRuntimeContext()
    import ray.data.block as rblock

    # block.py binds ``ray.get_runtime_context`` lazily through ``ray``.
    ray.get_runtime_context = rc.get_runtime_context
    ray.runtime_context.get_runtime_context = rc.get_runtime_context

    from ray.data._internal import size_estimator

    def _real_size(self, item):
        import ray.cloudpickle as cp

        return len(cp.dumps(item))

    size_estimator.SizeEstimator._real_size = _real_size

    from ray.data import aggregate as ragg
    from ray.data._internal import util as rutil
    from ray.data._internal.arrow_block import ArrowBlockAccessor
    from ray.data._internal.arrow_ops import transform_pyarrow
    from ray.data._internal.planner.exchange.aggregate_task_spec import (
        SortAggregateTaskSpec,
    )
    from ray.data._internal.planner.exchange.shuffle_task_spec import ShuffleTaskSpec
    from ray.data._internal.planner.exchange.sort_task_spec import (
        SortKey,
        SortTaskSpec,
    )
    from ray.data.block import BlockAccessor

    _loaded = types.SimpleNamespace(
        ray=ray,
        block=rblock,
        BlockAccessor=BlockAccessor,
        ArrowBlockAccessor=ArrowBlockAccessor,
        transform_pyarrow=transform_pyarrow,
        util=rutil,
        find_partitions=rutil.find_partitions,
        find_partition_index=rutil.find_partition_index,
        NULL_SENTINEL=rutil.NULL_SENTINEL,
        SortKey=SortKey,
        SortTaskSpec=SortTaskSpec,
        ShuffleTaskSpec=ShuffleTaskSpec,
        SortAggregateTaskSpec=SortAggregateTaskSpec,
        aggregate=ragg,
    )
    return _loaded
