"""Writes tests/golden/reference_hparams_defaults.json: every key of the reference's global HParams object with its
default, read from /root/reference/hparams.py:11-225 by parsing the keyword arguments of the HParams(...) call with
`ast` (TensorFlow is not importable here).  Run in the build container, where /root/reference exists."""
import ast
import json
import os

HERE = os.path.dirname(os.path.abspath(__file__))


def main(ref="/root/reference/hparams.py"):
    tree = ast.parse(open(ref).read())
    call = [n for n in ast.walk(tree) if isinstance(n, ast.Call) and getattr(n.func, "attr", "") == "HParams"][0]
    vals = {kw.arg: ast.literal_eval(kw.value) for kw in call.keywords}
    out = {k: (list(v) if isinstance(v, (list, tuple)) else v) for k, v in sorted(vals.items())}
    with open(os.path.join(HERE, "reference_hparams_defaults.json"), "w") as f:
        json.dump(out, f, indent=0)
    print(len(out), "keys")


if __name__ == "__main__":
    main()
