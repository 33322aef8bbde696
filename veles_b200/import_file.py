"""Import a workflow/config file as a module or as part of its package
(reference: veles/import_file.py:50-95)."""
import importlib
import importlib.util
import os
import sys


def get_file_package_and_module(file_name):
    package_dir = os.path.dirname(os.path.abspath(file_name))
    module_name = os.path.splitext(os.path.basename(file_name))[0]
    return os.path.basename(package_dir), module_name


def import_file_as_package(file_name):
    """Import ``file_name`` through its parent package (needs an __init__.py there)."""
    package_dir = os.path.dirname(os.path.abspath(file_name))
    if not os.path.exists(os.path.join(package_dir, "__init__.py")):
        raise ImportError("%s is not a package" % package_dir)
    package_name, module_name = get_file_package_and_module(file_name)
    parent = os.path.dirname(package_dir)
    if parent not in sys.path:
        sys.path.insert(0, parent)
    return importlib.import_module("%s.%s" % (package_name, module_name))


def import_file_as_module(file_name):
    file_name = os.path.abspath(file_name)
    if os.path.isdir(file_name):
        raise IsADirectoryError(file_name)
    if not os.path.exists(file_name):
        raise FileNotFoundError(file_name)
    _, module_name = get_file_package_and_module(file_name)
    spec = importlib.util.spec_from_file_location(module_name, file_name)
    module = importlib.util.module_from_spec(spec)
    sys.modules[module_name] = module
    sys.path.insert(0, os.path.dirname(file_name))
    try:
        spec.loader.exec_module(module)
    finally:
        sys.path.pop(0)
    return module


def try_to_import_file(file_name):
    try:
        return import_file_as_package(file_name)
    except Exception as e:  # noqa: BLE001
        try:
            return import_file_as_module(file_name)
        except Exception as e2:  # noqa: BLE001
            return e, e2
