"""Fixture loader for openbeam's YAML structure-definition files (docs/structure-definition-format.rst): reads the
sections the reference's parser reads (libopenbeam/src/CFiniteElementProblem_parse_file.cpp:163-228 — parameters,
beam_sections, nodes, elements, constraints, node_loads, element_loads) into a ``Structure``.

This is the caller side of the path ("later" row), not part of it: a loader for the example files and the
test fixtures, not the reference's parser — no error recovery, no warnings list, no labels.  Expressions are arithmetic
over the ``parameters`` section (the reference evaluates them with exprtk: parse_file.cpp:326-362); here a small AST
walker accepts numbers, names, + - * / ** (and ^), unary signs, parentheses and a few functions.
"""
from __future__ import annotations

import ast
import math
import operator

import yaml

from .structure import build_structure

_BIN = {ast.Add: operator.add, ast.Sub: operator.sub, ast.Mult: operator.mul, ast.Div: operator.truediv,
        ast.Pow: operator.pow, ast.Mod: operator.mod}
_FUN = {"sin": math.sin, "cos": math.cos, "tan": math.tan, "sqrt": math.sqrt, "abs": abs, "exp": math.exp, "log": math.log,
        "deg2rad": math.radians, "rad2deg": math.degrees}
_CONST = {"pi": math.pi, "e": math.e}


def evaluate(expr, params):
    """EvaluationContext::evaluate (types.h:161-166) for the arithmetic subset the example files use"""
    if isinstance(expr, (int, float)) and not isinstance(expr, bool):
        return float(expr)
    tree = ast.parse(str(expr).replace("^", "**").strip(), mode="eval")

    def ev(n):
        if isinstance(n, ast.Expression):
            return ev(n.body)
        if isinstance(n, ast.Constant) and isinstance(n.value, (int, float)):
            return float(n.value)
        if isinstance(n, ast.Name):
            if n.id in params:
                return params[n.id]
            if n.id in _CONST:
                return _CONST[n.id]
            raise ValueError("unknown parameter '%s' in expression '%s'" % (n.id, expr))
        if isinstance(n, ast.BinOp) and type(n.op) in _BIN:
            return _BIN[type(n.op)](ev(n.left), ev(n.right))
        if isinstance(n, ast.UnaryOp) and isinstance(n.op, (ast.USub, ast.UAdd)):
            v = ev(n.operand)
            return -v if isinstance(n.op, ast.USub) else v
        if isinstance(n, ast.Call) and isinstance(n.func, ast.Name) and n.func.id in _FUN and not n.keywords:
            return _FUN[n.func.id](*[ev(a) for a in n.args])
        raise ValueError("unsupported expression '%s'" % expr)

    return float(ev(tree))


def load_structure(path):
    """examples-structures/*.yaml -> Structure (the arrays include/obgpu.h takes)"""
    with open(path) as f:
        doc = yaml.safe_load(f)
    params = {}
    for k, v in (doc.get("parameters") or {}).items():   # parse_file.cpp:326-362: in file order, may use earlier ones
        params[str(k)] = evaluate(v, params)
    E = lambda x: evaluate(x, params)
    sections = {}
    for s in doc.get("beam_sections") or []:              # :365-420
        sections[s["name"]] = {k: v for k, v in s.items() if k != "name"}
    nodes = {}
    for n in doc["nodes"]:                                # :424-500; ids may come in any order, the list auto-grows
        i = int(round(E(n["id"])))
        c = n["coords"]
        nodes[i] = (E(c[0]), E(c[1])) + ((E(n["rot_z"]),) if "rot_z" in n else ())
    missing = [i for i in range(max(nodes) + 1) if i not in nodes]
    if missing:
        raise ValueError("Undefined node IDs: %s" % " ".join(map(str, missing)))
    node_list = [nodes[i] for i in range(len(nodes))]
    elements, esec = [], {}
    for k, e in enumerate(doc["elements"]):               # :514-600
        a, b = (int(round(E(x))) for x in e["nodes"])
        src = sections[e["section"]] if "section" in e else e
        t = e["type"]
        if t.startswith("BEAM2D"):
            par = (E(src["E"]), E(src["A"]), E(src.get("Iz", 0.0)) if t != "BEAM2D_AA" or "Iz" in src else 0.0)
        elif t in ("SPRING_1D", "SPRING_TORSION"):
            par = (E(src["K"]), 0.0, 0.0)
        elif t == "SPRING_XY":
            par = (E(src["Kx"]), E(src["Ky"]), 0.0)
        elif t == "SPRING_DXDYRZ":
            par = (E(src["Kx"]), E(src["Ky"]), E(src["KRz"]))
        else:
            raise ValueError("Unknown element type '%s'" % t)
        esec["e%d" % k] = par
        elements.append((t, a, b, "e%d" % k))
    cons = [(int(round(E(c["node"]))), c["dof"].upper()) + ((E(c["value"]),) if "value" in c else ())
            for c in doc.get("constraints") or []]        # :611-700
    nloads = [(int(round(E(l["node"]))), l["dof"].upper(), E(l["value"])) for l in doc.get("node_loads") or []]   # :708-780
    eloads = []
    for l in doc.get("element_loads") or []:              # :789-850 + each load's loadParamsFromSet
        el = int(round(E(l["element"])))
        t = l["type"].upper()
        d = (E(l.get("DX", 0)), E(l.get("DY", 0)), E(l.get("DZ", 0)))
        if t == "DISTRIB_UNIFORM":
            eloads.append((el, E(l["q"]), d))
        elif t in ("TRIANGULAR", "DISTRIB_TRIANGULAR"):
            eloads.append((el, "TRIANGULAR", E(l["q_ini"]), E(l["q_end"]), d))
        elif t == "CONCENTRATED":
            eloads.append((el, "CONCENTRATED", E(l["p"]), E(l["dist"]), d))
        elif t == "TEMPERATURE":
            eloads.append((el, "TEMPERATURE", E(l["deltaT"])))
        else:
            raise ValueError("Unknown element load type '%s'" % t)
    return build_structure(node_list, elements, esec, cons, nloads, eloads)
