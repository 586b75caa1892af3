"""``SortKey``: the sort specification object of the exchange operator surface.

Same constructor, accessors and error behaviour as the reference's
``SortKey`` (python/ray/data/_internal/planner/exchange/sort_task_spec.py:20-89),
so that code written against the reference (and its tests) reads unchanged.
"""
from __future__ import annotations

from typing import List, Optional, Tuple, Union


class SortKey:
    def __init__(
        self,
        key: Optional[Union[str, List[str]]] = None,
        descending: Union[bool, List[bool]] = False,
        boundaries: Optional[list] = None,
    ):
        if key is None:
            key = []
        elif isinstance(key, str):
            key = [key]
        if not isinstance(key, list) or any(not isinstance(k, str) for k in key):
            raise ValueError(f"Key must be a string or a list of strings, but got {key}.")
        if isinstance(descending, bool):
            descending = [descending] * len(key)
        elif isinstance(descending, list):
            if len(descending) != len(key):
                raise ValueError("Length of `descending` does not match the length of the key.")
        self._columns = key
        self._descending = descending
        if boundaries:
            if any(not isinstance(b, (int, float)) for b in boundaries):
                raise ValueError("The type of items in boundaries must be int or float.")
            boundaries = sorted(set(boundaries))
        self._boundaries = boundaries

    def get_columns(self) -> List[str]:
        return self._columns

    def get_descending(self) -> List[bool]:
        return self._descending

    def to_arrow_sort_args(self) -> List[Tuple[str, str]]:
        return [(c, "descending" if d else "ascending")
                for c, d in zip(self._columns, self._descending)]

    def to_pandas_sort_args(self) -> Tuple[List[str], List[bool]]:
        return self._columns, [not d for d in self._descending]

    def validate_schema(self, schema) -> None:
        """Raise ValueError if a key column is missing from ``schema``."""
        if schema is None:
            return
        if self._columns and len(schema.names) > 0:
            have = set(schema.names)
            for c in self._columns:
                if c not in have:
                    raise ValueError(
                        f"You specified the column '{c}', but there's no such column in the "
                        f"dataset. The dataset has columns: {schema.names}"
                    )

    @property
    def boundaries(self):
        return self._boundaries
