"""Metadata filter oracle. TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).

Evaluates, directly on a chunk's metadata dict, the filter documents that
``ContextManager._create_combined_filter`` builds (super_rag/context/context.py:74-111):

    {"or": [{"indexer": {"$in": index_types}}, {"indexer": {"$exists": False}}]}
    {"chat_id": chat_id}
    {"and": [<the two above>]}

The reference hands these to ``connector.search(filter=...)``, whose SeekDB connector drops them
(seekdb_connector.py:99-100), so there is no reference OUTPUT to pin against: parity unpinned; the semantics
restated here are the usual document-store ones the dicts spell out (``$in`` membership, ``$exists``,
bare value = equality, a missing field matches neither ``$in`` nor equality).
"""


def eval_filter(flt, meta) -> bool:
    if not flt:
        return True
    meta = meta or {}
    ok = True
    for key, cond in flt.items():
        if key == "and":
            ok = ok and all(eval_filter(c, meta) for c in cond)
        elif key == "or":
            ok = ok and any(eval_filter(c, meta) for c in cond)
        elif isinstance(cond, dict):
            present = meta.get(key) is not None
            for op, arg in cond.items():
                if op == "$in":
                    ok = ok and present and meta[key] in arg
                elif op == "$exists":
                    ok = ok and (present == bool(arg))
                else:
                    raise ValueError(f"unsupported operator {op!r}")
        else:
            ok = ok and meta.get(key) is not None and meta[key] == cond
    return ok


def filter_rows(flt, metas):
    """Boolean list: which rows a filter admits."""
    return [eval_filter(flt, m) for m in metas]
