"""Extracts the reference's click flags on the hot path into tests/golden/click_flags.json (build container only):
  utils/click_options.py:483-501  run_method_options   (4 flags)
  utils/click_options.py:503-542  approx_options       (13 parameters + --approx-output-dir)
and the flags scripts/image_net.sh / scripts/generated_scripts.py pass.  The option declarations are read from the
source with `ast` (importing the module would pull every model definition and a network download).

    python tests/golden/make_click_flags.py
"""
import ast
import json
import os
import re
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import ref_shim  # noqa: E402


def options_of(tree, func_name):
    out = []
    for node in ast.walk(tree):
        if isinstance(node, ast.FunctionDef) and node.name == func_name:
            for inner in node.body:
                if isinstance(inner, ast.FunctionDef):
                    for d in inner.decorator_list:
                        if isinstance(d, ast.Call) and getattr(d.func, "attr", "") == "option":
                            decl = ast.literal_eval(d.args[0])
                            kw = {}
                            for k in d.keywords:
                                if k.arg in ("default", "is_flag"):
                                    kw[k.arg] = ast.literal_eval(k.value)
                                elif k.arg == "type":
                                    kw["type"] = ast.unparse(k.value)
                            out.append(dict(decl=decl, **kw))
    return out


def main():
    root = ref_shim.REF_ROOT
    tree = ast.parse(open(os.path.join(root, "utils", "click_options.py")).read())
    flags = {"run_method_options": options_of(tree, "run_method_options"), "approx_options": options_of(tree, "approx_options")}
    used = set()
    for script in ("scripts/image_net.sh", "scripts/generated_scripts.py"):
        path = os.path.join(os.path.dirname(root) if not os.path.isdir(os.path.join(root, "scripts")) else root, script)
        if os.path.exists(path):
            used |= set(re.findall(r"(--[A-Za-z][A-Za-z0-9_\-]*)", open(path).read()))
    flags["script_flags"] = sorted(used)
    json.dump(flags, open(os.path.join(HERE, "click_flags.json"), "w"), indent=1, sort_keys=True)
    print({k: len(v) for k, v in flags.items()})


if __name__ == "__main__":
    main()
