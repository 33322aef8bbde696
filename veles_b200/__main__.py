"""Command line entry point: ``python -m veles_b200 [options] workflow.py [config.py] [root.x.y=z ...]``.

Parity target: ``veles/__main__.py:136-867``: argument parsing through the distributed registry,
logging setup, PRNG seeding (``_seed_random :483-537``: ``file:count[:dtype]``, hex string or ``-``),
workflow import as package else module (:396-424), config application — python via runpy, else
pickle, else json (:426-472), ``root.a.b=c`` overrides (:474-481), dispatch to regular /
``--optimize`` / ``--ensemble-*`` runs (:716-734), snapshot loading (:539-589), ``--dry-run``,
``--workflow-graph``, ``--dump-unit-attributes``, ``--result-file``; the workflow module must define
``run(load, main)`` (:799-818).
"""
from __future__ import annotations

import binascii
import errno
import json
import logging
import os
import pickle
import resource
import runpy
import sys

import numpy

from . import __version__, prng
from .cmdline import CommandLineBase
from .config import root, get
from .import_file import import_file_as_module, import_file_as_package
from .launcher import Launcher
from .logger import Logger
from .pickle2 import setup_pickle_debug
from .snapshotter import SnapshotterToFile
# make every argument-contributing class visible to the parser
from . import backends, accelerated_units, server, client, thread_pool  # noqa: F401
from .loader import base as _loader_base  # noqa: F401


class Main(Logger, CommandLineBase):
    EXIT_SUCCESS = 0
    EXIT_FAILURE = 1

    def __init__(self, interactive=False, *args, **kwargs):
        super(Main, self).__init__()
        self.interactive = interactive
        self._argv = None
        if interactive or args or kwargs:
            self._argv = Main.setup_argv(False, False, *args, **kwargs)
        self.launcher = None
        self.workflow = None
        self.load_called = False
        self.main_called = False
        self.args = None
        self.config_file = ""
        self.snapshot_file_name = None
        self.restored_from_snapshot = False

    # -- helpers --------------------------------------------------------------------------------------------
    @property
    def argv(self):
        return self._argv if self._argv is not None else sys.argv[1:]

    def _parse(self):
        parser = Main.init_parser(ignore_conflicts=True)
        self.args = parser.parse_args(self.argv)
        return self.args

    def _setup_logging(self):
        level = CommandLineBase.LOG_LEVEL_MAP[self.args.verbosity]
        Logger.setup_logging(level)
        self._install_sinks()
        for name in filter(None, self.args.debug.split(",")):
            logging.getLogger(name).setLevel(logging.DEBUG)
        if self.args.debug_pickle:
            setup_pickle_debug()

    def _install_sinks(self):
        """JSON-lines log / event files the web status searches (the reference's MongoDB logging:
        ``veles/logger.py:226-262``), when ``root.common.engine.log_sinks`` is set - by a site configuration, the
        configuration file or a ``root.common.engine.log_sinks=True`` override (hence called again after those)."""
        if getattr(self, "_sinks_installed", False) or not get(root.common.engine.log_sinks, False):
            return
        log_dir = get(root.common.dirs.logs, None)
        if not log_dir:
            return
        from .logger import events, install_log_sink
        stem = os.path.join(log_dir, "%s-%d" % (events.session, os.getpid()))
        install_log_sink(stem + ".logs.jsonl")
        events.open(stem + ".events.jsonl")
        self._sinks_installed = True

    def _rank_group_entropy(self, nbytes):
        """One process per GPU (WORLD_SIZE > 1): replicas must start from identical weights, so random seed material
        is drawn on rank 0 and broadcast (this also brings up the process group the data-parallel launcher uses)."""
        import torch
        import torch.distributed as dist
        if not dist.is_initialized():
            if torch.cuda.is_available() and getattr(self.args, "backend", "auto") not in ("numpy",):
                local = int(os.environ.get("LOCAL_RANK", "0"))
                torch.cuda.set_device(local % torch.cuda.device_count())
                dist.init_process_group("cpu:gloo,cuda:nccl")
            else:
                dist.init_process_group("gloo")
        buf = torch.from_numpy(numpy.frombuffer(os.urandom(nbytes), dtype=numpy.uint8).copy())
        dist.broadcast(buf, 0)
        return buf.numpy().tobytes()

    def _seed_random(self, rndvals):
        """Seed PRNG key 1, 2, ... from comma separated specs (file:count[:dtype] | hex | '-')."""
        world = int(os.environ.get("WORLD_SIZE", "1") or 1)
        replicated = world > 1 and getattr(self.args, "data_parallel", "auto") != "off" and \
            not self.args.master_address and not self.args.listen_address
        for i, rndval in enumerate(rndvals.split(",")):
            key = i + 1
            rg = prng.get(key)
            if replicated and (rndval == "-" or rndval.split(":")[0] in ("/dev/urandom", "/dev/random")):
                rg.seed(numpy.frombuffer(self._rank_group_entropy(64), dtype=numpy.uint32))
                continue
            if rndval == "-":
                rg.seed(numpy.frombuffer(os.urandom(64), dtype=numpy.uint32))
                continue
            if os.path.exists(rndval.split(":")[0]):
                parts = rndval.split(":")
                fname = parts[0]
                count = int(parts[1]) if len(parts) > 1 else 16
                dtype = numpy.dtype(parts[2]) if len(parts) > 2 else numpy.dtype(numpy.uint32)
                with open(fname, "rb") as f:
                    raw = f.read(count * dtype.itemsize)
                rg.seed(numpy.frombuffer(raw[:len(raw) // dtype.itemsize * dtype.itemsize], dtype=dtype))
                continue
            try:
                rg.seed(numpy.frombuffer(binascii.unhexlify(rndval), dtype=numpy.uint8))
            except (binascii.Error, ValueError):
                try:
                    rg.seed(int(rndval))
                except ValueError:
                    self.error("Cannot interpret the random seed %r", rndval)
                    sys.exit(Main.EXIT_FAILURE)

    def _load_model(self, fname_workflow):
        self.info("Loading workflow \"%s\"...", fname_workflow)
        try:
            return import_file_as_package(fname_workflow)
        except Exception as e:  # noqa: BLE001
            self.debug("Package import failed (%s), trying as a plain module", e)
        try:
            return import_file_as_module(fname_workflow)
        except FileNotFoundError:
            self.error("Workflow does not exist: \"%s\"", fname_workflow)
            sys.exit(errno.ENOENT)
        except IsADirectoryError:
            self.error("Workflow \"%s\" is a directory", fname_workflow)
            sys.exit(errno.EISDIR)

    def _apply_config(self, fname_config):
        if not fname_config:
            self.warning("Configuration path is empty")
            return
        self.info("Applying the configuration from %s...", fname_config)
        if not os.path.exists(fname_config):
            self.error("Configuration does not exist: \"%s\"", fname_config)
            sys.exit(errno.ENOENT)
        cfg = None
        if fname_config.endswith(".json"):
            with open(fname_config) as f:
                cfg = json.load(f)
        elif fname_config.endswith((".pickle", ".pkl")):
            with open(fname_config, "rb") as f:
                cfg = pickle.load(f)
        else:
            try:
                runpy.run_path(fname_config)
            except (SyntaxError, ValueError, UnicodeDecodeError):
                try:
                    with open(fname_config, "rb") as f:
                        cfg = pickle.load(f)
                except Exception:  # noqa: BLE001
                    with open(fname_config) as f:
                        cfg = json.load(f)
        if cfg is not None:
            for sub, val in cfg.items():
                getattr(root, sub).update(val)

    def _override_config(self, config_list):
        for line in config_list:
            if "=" not in line or not line.startswith("root."):
                self.error("Invalid configuration override %r (expected root.a.b=value)", line)
                sys.exit(Main.EXIT_FAILURE)
            try:
                exec(line, {"root": root, "numpy": numpy})  # noqa: S102 - same contract as the reference CLI
            except Exception:  # noqa: BLE001
                self.exception("Invalid configuration override %r", line)
                sys.exit(Main.EXIT_FAILURE)

    # -- the two callbacks handed to the workflow module's run() ----------------------------------------------------
    def _load(self, Workflow, **kwargs):
        self.load_called = True
        self.launcher = Launcher(self.interactive, **self._launcher_kwargs())
        if self.args.snapshot:
            self.workflow = self._load_workflow(self.args.snapshot)
            if self.workflow is not None:
                self.workflow.workflow = self.launcher
                self.restored_from_snapshot = True
                self.info("Resuming from the snapshot %s", self.args.snapshot)
        if self.workflow is None:
            self.workflow = Workflow(self.launcher, **kwargs)
        if self.args.result_file:
            self.workflow.result_file = self.args.result_file
        return self.workflow, self.restored_from_snapshot

    def _launcher_kwargs(self):
        a = self.args
        return dict(master_address=a.master_address, listen_address=a.listen_address, nodes=a.nodes,
                    test=a.test, result_file=a.result_file, stealth=a.stealth, log_file=a.log_file,
                    matplotlib_backend=a.matplotlib_backend, backend=getattr(a, "backend", None),
                    device=None)

    def _load_workflow(self, fname_snapshot):
        if fname_snapshot.startswith(("http://", "https://")):
            import tempfile
            import urllib.request
            tmp = os.path.join(tempfile.mkdtemp(prefix="veles_b200_"), os.path.basename(fname_snapshot))
            urllib.request.urlretrieve(fname_snapshot, tmp)
            fname_snapshot = tmp
        elif fname_snapshot.startswith("odbc://"):
            # odbc://<data source | sqlite:path>&<table>&<id>&<log id>[&<name>]   (reference __main__.py:543-571)
            parsed = fname_snapshot[len("odbc://"):].split("&")
            if len(parsed) not in (4, 5):
                self.error("Invalid ODBC source; template: odbc://<data source>&<table>&<id>&<log id>[&<name>] "
                           "(<table> and <log id> may be empty)")
                return None
            odbc, table, id_, log_id = parsed[:4]
            from .snapshotter import SnapshotterToDB
            try:
                return SnapshotterToDB.import_(odbc, table or "veles", id_, log_id or id_,
                                               parsed[4] if len(parsed) == 5 else None)
            except Exception as e:
                self.error("Failed to load the snapshot from the database: %s", e)
                return None
        try:
            return SnapshotterToFile.import_(fname_snapshot)
        except FileNotFoundError:
            self.error("Snapshot %s does not exist: starting from scratch", fname_snapshot)
            return None

    def _main(self, **kwargs):
        if self.args.dry_run == "load":
            return
        self.main_called = True
        from .accelerated_units import AcceleratedUnit
        AcceleratedUnit.apply_args(self.args)
        _loader_base.Loader.cli_train_ratio = getattr(self.args, "train_ratio", 1.0)
        kw = dict(kwargs)
        if getattr(self.args, "backend", None) and self.args.backend != "auto":
            kw.setdefault("backend", self.args.backend)
        if self.restored_from_snapshot:
            kw["snapshot"] = True
        self.launcher.args.device = getattr(self.args, "device", "")
        self.launcher.args.backend = getattr(self.args, "backend", "auto")
        self.launcher.initialize(**kw)
        if self.args.workflow_graph:
            if self.args.workflow_graph.lower().endswith(".svg"):       # a picture instead of the DOT text
                with open(self.args.workflow_graph, "w") as f:
                    f.write(self.workflow.generate_graph_svg())
                self.info("Saved the workflow graph picture to %s", self.args.workflow_graph)
            else:
                self.workflow.generate_graph(filename=self.args.workflow_graph, with_data_links=True)
        if self.args.dump_unit_attributes != "no":
            self._dump_unit_attributes(self.args.dump_unit_attributes == "all")
        if self.args.dry_run == "init" or self.args.visualize:
            return
        try:
            self.launcher.run()
        finally:
            self.launcher.stop()

    def _dump_unit_attributes(self, arrays=True):
        import pprint
        for unit in self.workflow.units:
            print("=" * 80)
            print(repr(unit))
            attrs = {}
            for k, v in sorted(unit.__dict__.items()):
                if k.endswith("_") or callable(v):
                    continue
                if not arrays and hasattr(v, "mem"):
                    v = "<Array %s>" % (getattr(v, "shape", ()),)
                attrs[k] = v
            pprint.pprint(attrs, width=120)

    # -- dispatch --------------------------------------------------------------------------------------------------------
    def _run_core(self, wm):
        from .genetics.config import fix_config
        if self.args.optimize != "no":
            from .genetics import optimization_workflow
            size, _, gens = self.args.optimize.partition(":")
            return optimization_workflow.run(self._load, self._main, model=wm, size=int(size),
                                             generations=int(gens) if gens else None,
                                             argv=self.argv, config_file=self.config_file)
        if self.args.ensemble_train:
            from .ensemble import model_workflow
            n, _, ratio = self.args.ensemble_train.partition(":")
            return model_workflow.run(self._load, self._main, model=wm, size=int(n),
                                      train_ratio=float(ratio) if ratio else 1.0, argv=self.argv,
                                      config_file=self.config_file, result_file=self.args.result_file)
        if self.args.ensemble_test:
            from .ensemble import test_workflow
            return test_workflow.run(self._load, self._main, model=wm, input_file=self.args.ensemble_test,
                                     argv=self.argv, config_file=self.config_file,
                                     result_file=self.args.result_file)
        fix_config(root)
        return self.run_module(wm)

    def run_module(self, wm):
        if not hasattr(wm, "run") or not callable(wm.run):
            self.error("The workflow module %s must define run(load, main)", getattr(wm, "__name__", wm))
            sys.exit(Main.EXIT_FAILURE)
        wm.run(self._load, self._main)
        if not self.load_called:
            self.warning("load() was not called by the workflow module's run()")
        if not self.main_called and self.args.dry_run not in ("load",):
            self.warning("main() was not called by the workflow module's run()")

    def run(self):
        from .compat import install_veles_alias
        install_veles_alias()   # user workflows / snapshots may name the reference's module paths (veles.*)
        args = self._parse()
        if args.version:
            print("veles_b200", __version__)
            return Main.EXIT_SUCCESS
        if args.frontend:
            # interactive composer (reference: veles/__main__.py:258-327): serve the option form locally, then run
            # with the command line the user posted
            from .scripts.generate_frontend import serve
            composed = serve([a for a in self.argv if a != "--frontend"])
            if not composed:
                print("Cancelled.")
                return Main.EXIT_SUCCESS
            print("Running with the following command line: %s" % " ".join(composed))
            self.argv = [a for a in composed if a != "--frontend"]
            args = self._parse()
        if args.html_help:
            from .scripts.generate_frontend import generate
            print(generate(self.argv))
            return Main.EXIT_SUCCESS
        if not args.workflow:
            Main.init_parser(ignore_conflicts=True).print_help()
            return Main.EXIT_FAILURE
        if args.workflow == "forge":
            from .forge.forge_client import main as forge_main
            return forge_main(self.argv[1:])
        self._setup_logging()
        if not args.no_logo and not self.interactive and sys.stdout.isatty():
            print(CommandLineBase.LOGO)
        if args.background:
            self._daemonize()
        self._seed_random(args.random_seed)
        wm = self._load_model(args.workflow)
        self.config_file = args.config
        if args.config == "-":
            base = os.path.splitext(args.workflow)[0]
            self.config_file = base + "_config.py" if os.path.exists(base + "_config.py") else ""
        if self.config_file:
            self._apply_config(self.config_file)
        self._override_config(args.config_list)
        self._install_sinks()
        if args.dump_config:
            root.print_()
            return Main.EXIT_SUCCESS
        self._run_core(wm)
        peak = resource.getrusage(resource.RUSAGE_SELF).ru_maxrss
        from .memory import Watcher
        self.info("Peak resident memory: %.1f MB; peak device arrays: %.1f MB", peak / 1024.0,
                  Watcher.max_mem_in_use / 2.0 ** 20)
        return Main.EXIT_SUCCESS

    def _daemonize(self):
        if os.fork() > 0:
            os._exit(0)
        os.setsid()
        if os.fork() > 0:
            os._exit(0)
        devnull = os.open(os.devnull, os.O_RDWR)
        for fd in (0, 1, 2):
            os.dup2(devnull, fd)


def __run__(*args, **kwargs):
    m = Main(bool(args or kwargs), *args, **kwargs)
    return m.run()


if __name__ == "__main__":
    sys.exit(__run__())
