"""`python -m kvedge_b200 template|lint <chart> [--set ..] [--set-string ..] [--set-file ..]`
-- the two Helm commands BASELINE.json's configs 1-2 name, through the GPU-backed engine."""
import argparse
import sys

from . import helm


def main(argv=None) -> int:
    ap = argparse.ArgumentParser(prog="kvedge_b200")
    sub = ap.add_subparsers(dest="cmd", required=True)
    for name in ("template", "lint"):
        p = sub.add_parser(name)
        if name == "template":
            p.add_argument("release", nargs="?", default=helm.DEFAULT_RELEASE)
        p.add_argument("chart")
        p.add_argument("--set", action="append", default=[])
        p.add_argument("--set-string", action="append", default=[])
        p.add_argument("--set-file", action="append", default=[])
        p.add_argument("--namespace", "-n", default="default")
    a = ap.parse_args(argv)
    try:
        if a.cmd == "template":
            sys.stdout.buffer.write(helm.template(a.chart, a.release, set=a.set, set_string=a.set_string,
                                                  set_file=a.set_file, namespace=a.namespace))
            return 0
        ok, msgs = helm.lint(a.chart, set=a.set, set_file=a.set_file)
        print(f"==> Linting {a.chart}")
        for m in msgs:
            print(m)
        print(f"\n1 chart(s) linted, {0 if ok else 1} chart(s) failed")
        return 0 if ok else 1
    except helm.KvError as e:
        print(f"Error: {e}", file=sys.stderr)
        return 1


if __name__ == "__main__":
    sys.exit(main())
