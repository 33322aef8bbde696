"""Distributed argparse: every class whose metaclass is ``CommandLineArgumentsRegistry`` may
contribute flags through a static ``init_parser(**kwargs)`` (reference: veles/cmdline.py:61-278)."""
from __future__ import annotations

import argparse
import sys


class CommandLineArgumentsRegistry(type):
    classes = []

    def __init__(cls, name, bases, clsdict):
        super(CommandLineArgumentsRegistry, cls).__init__(name, bases, clsdict)
        init_parser = clsdict.get("init_parser", None)
        if init_parser is None:
            return
        if not isinstance(init_parser, staticmethod):
            raise TypeError("init_parser must be a static method since the class has "
                            "CommandLineArgumentsRegistry metaclass")
        CommandLineArgumentsRegistry.classes.append(cls)


class SortingRawDescriptionHelpFormatter(argparse.RawDescriptionHelpFormatter):
    def add_arguments(self, actions):
        actions = sorted(actions, key=lambda a: a.option_strings[-1].lstrip("-") if a.option_strings else "")
        super(SortingRawDescriptionHelpFormatter, self).add_arguments(actions)


class CommandLineBase(object):
    """Builds the global parser (the CLI flag inventory of SURVEY §5.6)."""

    LOGO = r"""
 __     __   _             ____ ____   ___   ___
 \ \   / /__| | ___  ___  | __ )___ \ / _ \ / _ \   Blackwell-native
  \ \ / / _ \ |/ _ \/ __| |  _ \ __) | | | | | | |  dataflow deep learning
   \ V /  __/ |  __/\__ \ | |_) / __/| |_| | |_| |  (sm_100a)
    \_/ \___|_|\___||___/ |____/_____|\___/ \___/
"""
    DRY_RUN_CHOICES = ["load", "init", "exec", "no"]
    LOG_LEVEL_MAP = {"debug": 10, "info": 20, "warning": 30, "error": 40}
    SPECIAL_OPTS = ["--help", "--html-help", "--version", "--frontend", "--dump-config"]

    @staticmethod
    def init_parser(sphinx=False, ignore_conflicts=False):
        parser = argparse.ArgumentParser(
            prog="veles_b200", description=CommandLineBase.LOGO if not sphinx else "",
            formatter_class=SortingRawDescriptionHelpFormatter,
            conflict_handler="resolve" if ignore_conflicts else "error")
        for cls in CommandLineArgumentsRegistry.classes:
            parser = cls.init_parser(parser=parser)
        parser.add_argument("--no-logo", default=False, action="store_true", help="Do not print the logo.")
        parser.add_argument("--version", action="store_true", help="Print version number and exit.")
        parser.add_argument("--html-help", action="store_true", help="Open the documentation.")
        parser.add_argument("--frontend", action="store_true", help="Open the command line builder form.")
        parser.add_argument("-v", "--verbosity", type=str, default="info",
                            choices=sorted(CommandLineBase.LOG_LEVEL_MAP), help="Logging level.")
        parser.add_argument("--debug", type=str, default="",
                            help="Comma separated class names that log at debug level.")
        parser.add_argument("--debug-pickle", default=False, action="store_true",
                            help="Locate the attribute which fails to pickle.")
        parser.add_argument("-r", "--random-seed", type=str, default="/dev/urandom:16",
                            help="Seed: file:count[:dtype], a hex string, or '-'; comma separated per PRNG key.")
        parser.add_argument("-w", "--snapshot", default="",
                            help="Workflow snapshot to resume from (path, http(s):// or odbc://).")
        parser.add_argument("--dump-config", default=False, action="store_true",
                            help="Print the resulting configuration and exit.")
        parser.add_argument("--dry-run", default="no", choices=CommandLineBase.DRY_RUN_CHOICES,
                            help="load: stop after loading; init: after initialize; exec: run but do not wait.")
        parser.add_argument("--visualize", default=False, action="store_true",
                            help="Render the workflow graph and exit.")
        parser.add_argument("--optimize", type=str, default="no",
                            help="Genetic hyper-parameter search: population size[:generations].")
        parser.add_argument("--ensemble-train", type=str, default="",
                            help="Train an ensemble: <number of models>:<train ratio>.")
        parser.add_argument("--ensemble-test", type=str, default="",
                            help="Evaluate an ensemble: path to the --result-file of --ensemble-train.")
        parser.add_argument("--workflow-graph", type=str, default="", help="Save the workflow graph (DOT) to file.")
        parser.add_argument("--dump-unit-attributes", default="no", choices=["no", "pretty", "all"],
                            help="Print unit __dict__-s after initialization.")
        parser.add_argument("workflow", nargs="?", default="", help="Path to the workflow Python file.")
        parser.add_argument("config", nargs="?", default="-",
                            help="Path to the configuration file ('-' = <workflow>_config.py, '' = none).")
        parser.add_argument("config_list", nargs="*", default=[], metavar="root.x.y=value",
                            help="Configuration overrides.")
        parser.add_argument("-b", "--background", default=False, action="store_true", help="Daemonize.")
        return parser

    @staticmethod
    def setup_argv(sys_argv=True, reset_argv=False, *args, **kwargs):
        """Turn ``veles_b200(workflow, config, **kwargs)`` arguments into an argv list."""
        argv = list(args)
        for key, val in kwargs.items():
            flag = "--" + key.replace("_", "-") if len(key) > 1 else "-" + key
            if isinstance(val, bool):
                if val:
                    argv.insert(0, flag)
            elif val is not None:
                argv[0:0] = [flag, str(val)]
        if sys_argv:
            if reset_argv:
                sys.argv = [sys.argv[0]] + argv
            else:
                sys.argv.extend(argv)
        return argv
