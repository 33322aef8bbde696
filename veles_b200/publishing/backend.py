"""Publishing backends selected by name (reference: ``veles/publishing/*_backend.py``: markdown, pdf via weasyprint,
confluence).  Here: ``markdown``, ``html`` (self-contained, plots inlined as SVG), ``json``; ``pdf`` and
``confluence`` are recognised and raise a clear error (weasyprint / network are unavailable)."""
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


class PDFBackend(Backend):
    MAPPING = "pdf"

    def render(self, info):
        raise RuntimeError("The PDF backend needs weasyprint, which is not installed; use \"html\" or \"markdown\"")


class ConfluenceBackend(Backend):
    MAPPING = "confluence"

    def render(self, info):
        raise RuntimeError("The Confluence backend needs network access; use \"html\" or \"markdown\"")
