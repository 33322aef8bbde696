"""Optional site-wide configuration hook: define ``update(root)`` here or in ~/.veles_b200/site_config.py
(reference: veles/site_config.py)."""


def update(root):
    pass
