"""Publishing backends selected by name (reference: ``veles/publishing/*_backend.py``: markdown, pdf via weasyprint,
confluence).  Here: ``markdown``, ``html`` (self-contained, plots inlined as SVG), ``json``, ``ipynb``, ``latex``, ``pdf``
(weasyprint when installed, else a built-in text PDF writer) and ``confluence`` (XML-RPC client in
``confluence.py``)."""
from __future__ import annotations

import json
import os

from ..mapped_object_registry import MappedObjectsRegistry


class BackendRegistry(MappedObjectsRegistry):
    mapping = "backends"
    backends = {}


class Backend(object, metaclass=BackendRegistry):
    MAPPING = None

    def __init__(self, **kwargs):
        self.file = kwargs.get("file")
        self.directory = kwargs.get("directory", ".")

    def target(self, name, ext):
        if self.file:
            return self.file
        os.makedirs(self.directory, exist_ok=True)
        return os.path.join(self.directory, "%s_report.%s" % (name, ext))

    def render(self, info):
        raise NotImplementedError


class MarkdownBackend(Backend):
    MAPPING = "markdown"

    def text(self, info):
        out = ["# %s" % info["name"], "", info.get("description") or "", "",
               "* id: `%s`" % info.get("id"), "* mode: %s, device: %s" % (info.get("mode"), info.get("device")),
               "* run time: %.1f s, epochs: %s" % (info.get("run_time", 0), info.get("epochs")),
               "* random seeds: %s" % info.get("seeds"), "", "## Results", ""]
        for k, v in sorted(info.get("results", {}).items()):
            if k not in ("Confusion", "Labels", "models"):
                out.append("* **%s**: %s" % (k, v))
        out += ["", "## Unit run times", "", "| unit | seconds |", "|---|---|"]
        out += ["| %s | %.4f |" % (n, t) for n, t in info.get("unit_times", [])]
        out += ["", "## Configuration", "", "```", json.dumps(info.get("config", {}), indent=1, default=str), "```", ""]
        if info.get("plots"):
            out += ["## Plots", ""] + ["![%s](%s)" % (os.path.basename(p), p) for p in info["plots"]]
        out += ["", "## Workflow graph", "", "```dot", info.get("graph", ""), "```", ""]
        return "\n".join(out)

    def render(self, info):
        path = self.target(info["name"], "md")
        with open(path, "w") as f:
            f.write(self.text(info))
        return path


class HTMLBackend(MarkdownBackend):
    MAPPING = "html"

    def render(self, info):
        import html
        body = ["<html><head><meta charset='utf-8'><title>%s</title></head><body><pre>" % html.escape(info["name"]),
                html.escape(self.text(dict(info, plots=[]))), "</pre>"]
        for p in info.get("plots", []):
            if p and p.endswith(".svg") and os.path.exists(p):
                with open(p) as f:
                    body.append(f.read())
        body.append("</body></html>")
        path = self.target(info["name"], "html")
        with open(path, "w") as f:
            f.write("\n".join(body))
        return path


class JSONBackend(Backend):
    MAPPING = "json"

    def render(self, info):
        path = self.target(info["name"], "json")
        with open(path, "w") as f:
            json.dump(info, f, indent=1, default=str)
        return path


class IPythonNotebookBackend(MarkdownBackend):
    """Report as a Jupyter notebook (nbformat 4): the sections of the Markdown report as text cells, the results and
    unit timings as data a code cell loads into ``results`` / ``unit_times`` for further analysis (the reference ships
    an ``ipynb_template.ipynb`` for its jinja2 backend: ``veles/publishing/jinja2_template_backend.py``)."""
    MAPPING = "ipynb"

    def render(self, info):
        def md(text):
            return {"cell_type": "markdown", "metadata": {}, "source": text.splitlines(True)}

        def code(text):
            return {"cell_type": "code", "metadata": {}, "execution_count": None, "outputs": [],
                    "source": text.splitlines(True)}

        sections = self.text(dict(info, plots=[])).split("\n## ")
        cells = [md(sections[0])] + [md("## " + sec) for sec in sections[1:]]
        plain = {k: v for k, v in info.get("results", {}).items() if k not in ("Confusion", "Labels", "models")}
        cells.append(code("import json\nresults = json.loads(%r)\nunit_times = json.loads(%r)\nresults\n"
                          % (json.dumps(plain, default=str), json.dumps(info.get("unit_times", []), default=str))))
        for p in info.get("plots", []):
            if p:
                cells.append(md("![%s](%s)" % (os.path.basename(p), p)))
        nb = {"cells": cells, "metadata": {"kernelspec": {"display_name": "Python 3", "language": "python",
                                                          "name": "python3"},
                                           "language_info": {"name": "python"}, "title": str(info["name"])},
              "nbformat": 4, "nbformat_minor": 5}
        path = self.target(info["name"], "ipynb")
        with open(path, "w") as f:
            json.dump(nb, f, indent=1)
        return path


class LatexBackend(Backend):
    """Report as a LaTeX article (``latex_template.tex`` of the reference's jinja2 backend): description, result and
    unit-time tables, the configuration verbatim, plots as figures when they are PDF / PNG files."""
    MAPPING = "latex"
    SPECIAL = {"\\": r"\textbackslash{}", "&": r"\&", "%": r"\%", "$": r"\$", "#": r"\#", "_": r"\_", "{": r"\{",
               "}": r"\}", "~": r"\textasciitilde{}", "^": r"\textasciicircum{}"}

    @classmethod
    def esc(cls, text):
        return "".join(cls.SPECIAL.get(c, c) for c in str(text))

    def text(self, info):
        e = self.esc
        out = [r"\documentclass[a4paper,10pt]{article}", r"\usepackage[utf8]{inputenc}", r"\usepackage{graphicx}",
               r"\usepackage{longtable}", r"\usepackage[margin=2cm]{geometry}",
               r"\title{%s}" % e(info["name"]), r"\date{}", r"\begin{document}", r"\maketitle",
               e(info.get("description") or ""), "",
               r"\begin{itemize}", r"\item id: \texttt{%s}" % e(info.get("id")),
               r"\item mode: %s, device: %s" % (e(info.get("mode")), e(info.get("device"))),
               r"\item run time: %.1f s, epochs: %s" % (info.get("run_time", 0), e(info.get("epochs"))),
               r"\item random seeds: %s" % e(info.get("seeds")), r"\end{itemize}",
               r"\section{Results}", r"\begin{longtable}{ll}"]
        for k, v in sorted(info.get("results", {}).items()):
            if k not in ("Confusion", "Labels", "models"):
                out.append(r"%s & %s \\" % (e(k), e(v)))
        out += [r"\end{longtable}", r"\section{Unit run times}", r"\begin{longtable}{lr}", r"unit & seconds \\ \hline"]
        out += [r"%s & %.4f \\" % (e(n), t) for n, t in info.get("unit_times", [])]
        out += [r"\end{longtable}", r"\section{Configuration}", r"\begin{verbatim}",
                json.dumps(info.get("config", {}), indent=1, default=str), r"\end{verbatim}"]
        figs = [p for p in info.get("plots", []) if p and p.lower().endswith((".pdf", ".png", ".jpg"))]
        if figs:
            out.append(r"\section{Plots}")
            for p in figs:
                out += [r"\begin{figure}[h]\centering", r"\includegraphics[width=0.8\textwidth]{%s}" % p,
                        r"\caption{%s}" % e(os.path.basename(p)), r"\end{figure}"]
        out += [r"\end{document}", ""]
        return "\n".join(out)

    def render(self, info):
        path = self.target(info["name"], "tex")
        with open(path, "w") as f:
            f.write(self.text(info))
        return path


def write_text_pdf(path, text, title="report", font_size=9, page=(595, 842), margin=40):
    """A self-contained PDF 1.4 writer for monospaced text reports (Courier, Latin-1): pages, xref, trailer.
    The reference renders HTML through weasyprint (``publishing/pdf_backend.py:48-90``), which is not available
    here; a report that is text plus tables does not need a layout engine."""
    leading = font_size * 1.25
    cols = int((page[0] - 2 * margin) / (font_size * 0.6))
    rows = int((page[1] - 2 * margin) / leading)
    lines = []
    for raw in text.replace("\t", "    ").split("\n"):
        if not raw:
            lines.append("")
        while raw:
            lines.append(raw[:cols])
            raw = raw[cols:]
    pages = [lines[k:k + rows] for k in range(0, max(len(lines), 1), rows)]

    def esc(t):
        t = t.encode("latin-1", "replace").decode("latin-1")
        return t.replace("\\", "\\\\").replace("(", "\\(").replace(")", "\\)")

    objs = []                                   # 1-based object bodies (bytes)

    def add(body):
        objs.append(body if isinstance(body, bytes) else body.encode("latin-1"))
        return len(objs)

    catalog = add("")                           # filled below
    pages_obj = add("")
    font = add("<< /Type /Font /Subtype /Type1 /BaseFont /Courier /Encoding /WinAnsiEncoding >>")
    kids = []
    for chunk in pages:
        ops = ["BT", "/F1 %d Tf" % font_size, "%g TL" % leading, "%d %g Td" % (margin, page[1] - margin - font_size)]
        for ln in chunk:
            ops.append("(%s) Tj T*" % esc(ln))
        ops.append("ET")
        stream = "\n".join(ops).encode("latin-1")
        content = add(b"<< /Length %d >>\nstream\n" % len(stream) + stream + b"\nendstream")
        kids.append(add("<< /Type /Page /Parent %d 0 R /MediaBox [0 0 %d %d] /Contents %d 0 R "
                        "/Resources << /Font << /F1 %d 0 R >> >> >>" % (pages_obj, page[0], page[1], content, font)))
    objs[catalog - 1] = ("<< /Type /Catalog /Pages %d 0 R >>" % pages_obj).encode("latin-1")
    objs[pages_obj - 1] = ("<< /Type /Pages /Count %d /Kids [%s] >>"
                           % (len(kids), " ".join("%d 0 R" % k for k in kids))).encode("latin-1")
    info = add("<< /Title (%s) /Producer (veles_b200) >>" % esc(title))
    out = bytearray(b"%PDF-1.4\n%\xe2\xe3\xcf\xd3\n")
    offsets = []
    for n, body in enumerate(objs, 1):
        offsets.append(len(out))
        out += b"%d 0 obj\n" % n + body + b"\nendobj\n"
    xref = len(out)
    out += b"xref\n0 %d\n0000000000 65535 f \n" % (len(objs) + 1)
    for off in offsets:
        out += b"%010d 00000 n \n" % off
    out += b"trailer\n<< /Size %d /Root %d 0 R /Info %d 0 R >>\nstartxref\n%d\n%%%%EOF\n" % (
        len(objs) + 1, catalog, info, xref)
    with open(path, "wb") as f:
        f.write(bytes(out))
    return len(pages)


class PDFBackend(MarkdownBackend):
    """Report as PDF: through weasyprint when it is installed (the reference's route), else as a paginated text PDF
    written by ``write_text_pdf``."""
    MAPPING = "pdf"

    def render(self, info):
        path = self.target(info["name"], "pdf")
        try:
            import weasyprint
        except ImportError:
            weasyprint = None
        if weasyprint is not None:
            import html
            doc = "<html><body><pre>%s</pre></body></html>" % html.escape(self.text(info))
            weasyprint.HTML(string=doc).write_pdf(path)
        else:
            self.pages = write_text_pdf(path, self.text(dict(info, graph="")), title=info["name"])
        return path


class ConfluenceBackend(MarkdownBackend):
    """Publishes the report as a Confluence page and attaches the plots (reference:
    ``publishing/confluence_backend.py:42-80``).  kwargs: ``server``, ``username``, ``password``, ``space``,
    optional ``page`` (default: the workflow name, made unique with " (n)") and ``parent``."""
    MAPPING = "confluence"

    def __init__(self, **kwargs):
        super(ConfluenceBackend, self).__init__(**kwargs)
        missing = [k for k in ("server", "username", "password", "space") if k not in kwargs]
        if missing:
            raise ValueError("ConfluenceBackend: missing %s" % ", ".join(missing))
        self.access = {k: kwargs[k] for k in ("server", "username", "password")}
        self.space = kwargs["space"]
        self.page = kwargs.get("page")
        self.parent = kwargs.get("parent")
        self.url = None

    def storage_xml(self, info):
        import html
        parts = ["<h1>%s</h1>" % html.escape(str(info["name"])), "<p>%s</p>" % html.escape(info.get("description") or ""),
                 "<ul>"]
        for k in ("id", "mode", "device", "epochs", "seeds"):
            parts.append("<li><strong>%s</strong>: %s</li>" % (k, html.escape(str(info.get(k)))))
        parts.append("<li><strong>run time</strong>: %.1f s</li></ul><h2>Results</h2><table><tbody>"
                     % info.get("run_time", 0))
        for k, v in sorted(info.get("results", {}).items()):
            if k not in ("Confusion", "Labels", "models"):
                parts.append("<tr><th>%s</th><td>%s</td></tr>" % (html.escape(str(k)), html.escape(str(v))))
        parts.append("</tbody></table><h2>Unit run times</h2><table><tbody>")
        for n, t in info.get("unit_times", []):
            parts.append("<tr><td>%s</td><td>%.4f</td></tr>" % (html.escape(str(n)), t))
        parts.append("</tbody></table><h2>Configuration</h2>")
        parts.append('<ac:structured-macro ac:name="code"><ac:plain-text-body><![CDATA[%s]]></ac:plain-text-body>'
                     '</ac:structured-macro>' % json.dumps(info.get("config", {}), indent=1, default=str)
                     .replace("]]>", "]] >"))
        for p in info.get("plots", []):
            parts.append('<p><ac:image><ri:attachment ri:filename="%s" /></ac:image></p>'
                         % html.escape(os.path.basename(p)))
        return "".join(parts)

    def render(self, info):
        from .confluence import Confluence
        conf = Confluence(**self.access)
        page = self.page
        if page is None:
            page, index = str(info["name"]), 1
            while conf.get_page_summary(page, self.space) is not None:
                page = "%s (%d)" % (info["name"], index)
                index += 1
        published = conf.store_page_content(page, self.space, self.storage_xml(info), convert_wiki=False,
                                            parent=self.parent)
        files = {os.path.basename(p): p for p in info.get("plots", []) if p and os.path.exists(p)}
        if info.get("graph"):
            files["workflow.dot"] = info["graph"].encode("utf-8")
        if files:
            conf.attach_file(page, self.space, files)
        self.page = page
        self.url = published.get("url")
        return self.url or page
