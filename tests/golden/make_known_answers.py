#!/usr/bin/env python
"""Extract the known-answer vectors of the reference's own pytest files into a fixture.

Run in the build container only (the reference is mounted at /root/reference):
    python tests/golden/make_known_answers.py
Reads /root/reference/tests/test_observation.py and test_mask.py with ``ast`` (nothing is
imported or executed: the reference needs a JVM) and writes tests/golden/known_answers.json:

  obs.cells   name -> 29-vector, the literals of test_observation.py:22-57,98-105 with the
              p2_* variants derived as the test does (:46-49, owner planes 10:13 = [0,0,1])
  obs.checks  [map, env, y, x, cell name] for every assert_array_equal (:60-78,108)
  obs.total   3072, the asserted sum over both views (:84)
  mask.rows   [cell index, 78-vector] of test_mask.py:28-84 (map 4x4/baseTwoWorkers4x4, 1 bot env vs passiveAI)
"""
import ast
import json
import os

REF = "/root/reference/tests"
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "known_answers.json")


def first_list(node):
    for n in ast.walk(node):
        if isinstance(n, ast.List) and n.elts and all(isinstance(e, (ast.Constant, ast.UnaryOp)) for e in n.elts):
            return [int(ast.literal_eval(e)) for e in n.elts]
    return None


def main():
    # ---------------- observation
    tree = ast.parse(open(os.path.join(REF, "test_observation.py")).read())
    cells, checks, maps = {}, [], []
    fn = next(n for n in tree.body if isinstance(n, ast.FunctionDef) and n.name == "test_observation")
    for node in ast.walk(fn):
        if isinstance(node, ast.keyword) and node.arg == "map_paths":
            maps.append(ast.literal_eval(node.value)[0])
    cur_map = -1
    for stmt in fn.body:
        src = ast.unparse(stmt)
        if isinstance(stmt, ast.Assign) and "MicroRTSGridModeVecEnv" in src:
            cur_map += 1
        elif isinstance(stmt, ast.Assign) and isinstance(stmt.targets[0], ast.Name):
            name = stmt.targets[0].id
            v = first_list(stmt.value)
            if v is not None and len(v) == 29:
                cells[name] = v
            elif src.endswith(".copy()"):
                cells[name] = list(cells[stmt.value.func.value.id])
        elif isinstance(stmt, ast.Assign) and isinstance(stmt.targets[0], ast.Subscript):
            tgt = stmt.targets[0]
            sl = tgt.slice
            lo, hi = ast.literal_eval(sl.lower), ast.literal_eval(sl.upper)
            cells[tgt.value.id][lo:hi] = first_list(stmt.value)
        elif isinstance(stmt, ast.Expr) and "assert_array_equal" in src:
            a, b = stmt.value.args
            idx = []
            n = a
            while isinstance(n, ast.Subscript):
                idx.append(ast.literal_eval(n.slice))
                n = n.value
            x, y, env = idx  # next_obs[env][y][x]
            checks.append([maps[cur_map], env, y, x, b.id])
    total = None
    for node in ast.walk(fn):
        if isinstance(node, ast.Assert) and "3072" in ast.unparse(node):
            total = 3072
    assert len(checks) == 17 and total == 3072 and set(c[4] for c in checks) <= set(cells)

    # ---------------- mask
    tree = ast.parse(open(os.path.join(REF, "test_mask.py")).read())
    rows = []
    for node in ast.walk(tree):
        if isinstance(node, ast.Call) and ast.unparse(node.func).endswith("assert_array_equal"):
            a, b = node.args
            cell = ast.literal_eval(a.slice)[1]
            v = first_list(b)
            assert len(v) == 78
            rows.append([cell, v])
    assert [r[0] for r in rows] == [1, 4, 5]
    json.dump({"obs": {"cells": cells, "checks": checks, "total": total}, "mask": {"map": "maps/4x4/baseTwoWorkers4x4.xml", "rows": rows}}, open(OUT, "w"))
    print("wrote", OUT, len(cells), "cells", len(checks), "checks", len(rows), "mask rows")


if __name__ == "__main__":
    main()
