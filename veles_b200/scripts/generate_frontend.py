"""Generate a static HTML form listing every command line option (the reference's ``--frontend`` builder:
``veles/scripts/generate_frontend.py``): ``python -m veles_b200.scripts.generate_frontend out.html``."""
from __future__ import annotations

import html
import sys


def generate():
    from ..__main__ import Main
    parser = Main.init_parser(ignore_conflicts=True)
    rows = []
    for act in parser._actions:
        if not act.option_strings:
            continue
        rows.append("<tr><td><code>%s</code></td><td>%s</td><td>%s</td></tr>"
                    % (html.escape(", ".join(act.option_strings)), html.escape(str(act.default)),
                       html.escape(act.help or "")))
    return ("<html><head><meta charset='utf-8'><title>veles_b200 command line</title></head><body>"
            "<h2>veles_b200 [options] workflow.py [config.py] [root.x.y=value ...]</h2>"
            "<table border=1><tr><th>option</th><th>default</th><th>meaning</th></tr>%s</table></body></html>"
            % "\n".join(rows))


def main(argv=None):
    argv = argv if argv is not None else sys.argv[1:]
    text = generate()
    if argv:
        with open(argv[0], "w") as f:
            f.write(text)
    else:
        print(text)
    return 0


if __name__ == "__main__":
    sys.exit(main())
