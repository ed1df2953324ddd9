"""Build-container only: freeze the list of names the reference's callers import from `ProposalNetwork.*`.

    python tests/golden/make_mirror_imports.py   ->   tests/golden/mirror_imports.json

Parses (ast, no execution) every .py under /root/reference/cubercnn and /root/reference/tools plus the on-path modules
of ProposalNetwork itself (proposals/proposals.py, scoring/scorefunction.py) for `import ProposalNetwork.x.y [as z]` and
`from ProposalNetwork.x.y import a, b`.  tests/test_mirror_imports.py asserts every entry resolves from 3dod_b200/."""
import ast
import json
import os

REF = "/root/reference"
ROOTS = ["cubercnn", "tools", "ProposalNetwork/proposals/proposals.py", "ProposalNetwork/scoring/scorefunction.py"]
# modules of the reference that are explicitly off the path (SURVEY.md 8(a) "not on the path"): importing them is not
# part of the drop-in contract
OFF_PATH = {"ProposalNetwork.scoring.convex_outline"}


def files():
    for r in ROOTS:
        p = os.path.join(REF, r)
        if os.path.isfile(p):
            yield p
            continue
        for d, _, fs in os.walk(p):
            for f in fs:
                if f.endswith(".py"):
                    yield os.path.join(d, f)


def main():
    rows = []
    for path in sorted(files()):
        try:
            tree = ast.parse(open(path).read())
        except SyntaxError:
            continue
        rel = os.path.relpath(path, REF)
        for node in ast.walk(tree):
            if isinstance(node, ast.ImportFrom) and node.module and node.module.startswith("ProposalNetwork"):
                if node.module in OFF_PATH:
                    continue
                rows.append({"where": "%s:%d" % (rel, node.lineno), "module": node.module,
                             "names": [a.name for a in node.names]})
            elif isinstance(node, ast.Import):
                for a in node.names:
                    if a.name.startswith("ProposalNetwork") and a.name not in OFF_PATH:
                        rows.append({"where": "%s:%d" % (rel, node.lineno), "module": a.name, "names": []})
    out = os.path.join(os.path.dirname(os.path.abspath(__file__)), "mirror_imports.json")
    json.dump(rows, open(out, "w"), indent=1)
    print(out, len(rows), "import statements")


if __name__ == "__main__":
    main()
