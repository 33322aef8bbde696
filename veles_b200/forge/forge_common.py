"""Package manifest helpers shared by the forge client and server (reference: veles/forge/forge_common.py)."""
import io
import json
import os
import tarfile

REQUIRED_MANIFEST_FIELDS = {"name", "workflow", "configuration", "short_description", "long_description", "author"}


def validate_manifest(manifest):
    missing = REQUIRED_MANIFEST_FIELDS - set(manifest)
    if missing:
        raise ValueError("manifest.json lacks the fields %s" % sorted(missing))
    return manifest


def pack(directory):
    """tar.gz of a model directory (must contain manifest.json)."""
    with open(os.path.join(directory, "manifest.json")) as f:
        manifest = validate_manifest(json.load(f))
    bio = io.BytesIO()
    with tarfile.open(fileobj=bio, mode="w:gz") as tar:
        for base, _, names in os.walk(directory):
            for n in sorted(names):
                full = os.path.join(base, n)
                tar.add(full, arcname=os.path.relpath(full, directory))
    return manifest, bio.getvalue()


def unpack(blob, directory):
    os.makedirs(directory, exist_ok=True)
    with tarfile.open(fileobj=io.BytesIO(blob), mode="r:gz") as tar:
        tar.extractall(directory, filter="data")
    with open(os.path.join(directory, "manifest.json")) as f:
        return json.load(f)
