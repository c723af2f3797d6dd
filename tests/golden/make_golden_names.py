"""Extract the Keras registration names of the reference's layer classes (data only, via ast; the reference is not
imported because TensorFlow is absent).  Output: tests/golden/registered_names.json  {class name: 'package>name'}.

    python tests/golden/make_golden_names.py /root/reference
"""
import ast
import json
import os
import sys


def main(ref):
    out = {}
    for base in ("kgcnn/layers", "kgcnn/model", "kgcnn/literature"):
        for root, _, files in os.walk(os.path.join(ref, base)):
            for f in sorted(files):
                if not f.endswith(".py"):
                    continue
                tree = ast.parse(open(os.path.join(root, f)).read())
                for node in ast.walk(tree):
                    if not isinstance(node, ast.ClassDef):
                        continue
                    for dec in node.decorator_list:
                        if isinstance(dec, ast.Call) and getattr(dec.func, "attr", "") == "register_keras_serializable":
                            kw = {k.arg: ast.literal_eval(k.value) for k in dec.keywords}
                            out[node.name] = "%s>%s" % (kw.get("package", "Custom"), kw.get("name", node.name))
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "registered_names.json")
    with open(path, "w") as fh:
        json.dump(dict(sorted(out.items())), fh, indent=0)
    print(len(out), "classes ->", path)


if __name__ == "__main__":
    main(sys.argv[1] if len(sys.argv) > 1 else "/root/reference")
