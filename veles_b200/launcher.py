"""Launcher: owns the execution mode, the device, the thread pool and the job-control agent.

Parity target: ``veles/launcher.py:100-906``: mode = master if ``-l``, slave if ``-m``, else
standalone (:332-356); creates ``Server`` / ``Client`` (:476-491), the ``Device`` (not on the
master, :493-494), initializes the workflow with ``device=`` (:510-517), runs it and blocks
(:550-571); status heartbeats to the web status server (:852-885); ``filter_argv`` (:75-96)
and SSH/local spawning of slave nodes (:808-842).

Differences: no Twisted reactor (a plain event + the thread pool), slaves on the local box
are spawned with ``subprocess`` (one per GPU: ``CUDA_VISIBLE_DEVICES=i``) instead of SSH unless
a remote host is named, plotting is headless.  Gradient exchange for data-parallel training
does NOT go through master/slave any more (veles_b200.parallel); this control plane carries
genetics / ensemble jobs and the reference's job protocol for compatibility.
"""
from __future__ import annotations

import argparse
import json
import os
import socket
import subprocess
import sys
import threading
import time
import uuid

from .cmdline import CommandLineArgumentsRegistry
from .config import root, get
from .logger import Logger
from .thread_pool import ThreadPool


def filter_argv(argv, *blacklist):
    """Remove ``blacklist`` options (with their values) from an argv list.
    Entries ending with "=" or listed in VALUE_OPTS take a value."""
    out = []
    skip = False
    for i, arg in enumerate(argv):
        if skip:
            skip = False
            continue
        matched = False
        for b in blacklist:
            takes_value = b.endswith("=")
            name = b.rstrip("=")
            if arg == name:
                matched = True
                skip = takes_value
                break
            if arg.startswith(name + "=") or (len(name) == 2 and arg.startswith(name) and takes_value and len(arg) > 2):
                matched = True
                break
        if not matched:
            out.append(arg)
    return out


class Launcher(Logger, metaclass=CommandLineArgumentsRegistry):
    """Workflow launcher and mode selector."""

    def __init__(self, interactive=False, **kwargs):
        super(Launcher, self).__init__()
        parser = Launcher.init_parser(parser=argparse.ArgumentParser(add_help=False))
        # interactive (in-process / notebook / test) launchers take their options from kwargs only
        self.args, _ = parser.parse_known_args(kwargs.get("argv", [] if interactive else sys.argv[1:]))
        for k, v in kwargs.items():
            if hasattr(self.args, k):
                setattr(self.args, k, v)
        self.interactive = interactive
        self.id = str(uuid.uuid4()) if not self.args.master_address else None
        self.log_id = self.args.log_id or self.id
        self.workflow = None
        self.device = kwargs.get("device")
        self.agent = None
        self.dp_ = None
        self._thread_pool = None
        self._initialized = False
        self._running = False
        self._finished = threading.Event()
        self._start_time = None
        self.stopped = False
        self.graphics = None
        self._notify_thread = None
        self._children = []
        self.testing = bool(self.args.test)
        self.result_file = self.args.result_file

    @staticmethod
    def init_parser(**kwargs):
        parser = kwargs.get("parser", argparse.ArgumentParser())
        parser.add_argument("-m", "--master-address", type=str, default=kwargs.get("master_address", ""),
                            help="Slave mode: master's host:port.").mode = ("slave",)
        parser.add_argument("-l", "--listen-address", type=str, default=kwargs.get("listen_address", ""),
                            help="Master mode: host:port to listen on.").mode = ("master",)
        parser.add_argument("-n", "--nodes", type=str, default="",
                            help="Slave nodes to launch: host/backend:devs[xN], comma separated "
                                 "(\"localhost/b200:0-7\").").mode = ("master",)
        parser.add_argument("--max-nodes", type=int, default=0, help="Limit on the number of slaves.")
        parser.add_argument("--slave-launch-transform", type=str, default="",
                            help="printf-style wrapper (%%s = the command) applied to slave command lines.")
        parser.add_argument("--data-parallel", type=str, default="auto", choices=["auto", "off", "fused", "nccl", "host"],
                            help="With WORLD_SIZE > 1 (one process per GPU): synchronous data parallelism through "
                                 "the fused all-reduce+update kernel (auto on GPUs), NCCL, or host all-reduce (numpy).")
        parser.add_argument("--yarn-nodes", type=str, default="",
                            help="Discover the slave nodes from this YARN ResourceManager (host[:8088]): every "
                                 "RUNNING node of /ws/v1/cluster/nodes is launched as host/auto:0.")
        parser.add_argument("-t", "--test", default=False, action="store_true",
                            help="Evaluation mode: run the workflow on the test data without training.")
        parser.add_argument("-p", "--matplotlib-backend", type=str, nargs="?", const="", default="Agg",
                            help="Matplotlib backend ('' disables plotting); GUI backends are unavailable here.")
        parser.add_argument("--no-graphics-client", default=False, action="store_true",
                            help="Do not start the plot renderer.")
        parser.add_argument("--graphics-pickle-debug", default=False, action="store_true")
        parser.add_argument("--pdb-on-finish", default=False, action="store_true",
                            help="Drop into pdb after the workflow finishes.")
        parser.add_argument("-s", "--stealth", default=False, action="store_true",
                            help="Do not report status to the web status server.")
        parser.add_argument("-f", "--log-file", type=str, default="", help="Redirect logs to this file.")
        parser.add_argument("--log-file-pid", default=False, action="store_true",
                            help="Append the pid to the log file name.")
        parser.add_argument("-g", "--log-mongo", type=str, nargs="?", const="127.0.0.1:27017", default="",
                            help="Duplicate logs to MongoDB (needs pymongo).")
        parser.add_argument("-i", "--log-id", type=str, default="", help="Log session identifier.")
        parser.add_argument("--result-file", type=str, default="", help="Write the results JSON here.")
        parser.add_argument("--manhole", default=False, action="store_true",
                            help="Run the embedded interactive shell on the UNIX socket /tmp/manhole-<pid> "
                                 "(also opened on demand by SIGUSR2); attach with python -m veles_b200.manhole <pid>.")
        return parser

    # -- mode -------------------------------------------------------------------------------------------------
    @property
    def mode(self):
        if self.args.listen_address:
            return "master"
        if self.args.master_address:
            return "slave"
        return "standalone"

    @property
    def is_master(self):
        return self.mode == "master"

    @property
    def is_slave(self):
        return self.mode == "slave"

    @property
    def is_standalone(self):
        return self.mode == "standalone"

    @property
    def is_main(self):
        return True

    @property
    def is_initialized(self):
        return self._initialized

    @property
    def is_running(self):
        return self._running

    @property
    def thread_pool(self):
        if self._thread_pool is None:
            tp = root.common.engine.thread_pool
            self._thread_pool = ThreadPool(minthreads=get(tp.minthreads, 2), maxthreads=get(tp.maxthreads, 2),
                                           name="launcher")
        return self._thread_pool

    @property
    def plots_endpoints(self):
        return getattr(self.graphics, "endpoints", []) if self.graphics is not None else []

    @property
    def start_time(self):
        return self._start_time

    # -- container protocol for the workflow ----------------------------------------------------------------------
    def add_ref(self, workflow):
        self.workflow = workflow
        if self.result_file and hasattr(workflow, "result_file"):
            workflow.result_file = self.result_file

    def del_ref(self, workflow):
        if self.workflow is workflow:
            self.workflow = None

    def on_workflow_finished(self):
        self._finished.set()

    # -- life cycle ---------------------------------------------------------------------------------------------------
    def initialize(self, **kwargs):
        if self.workflow is None:
            raise RuntimeError("Launcher.initialize(): no workflow was attached (Workflow(launcher, ...))")
        if self.args.log_file:
            name = self.args.log_file + (".%d" % os.getpid() if self.args.log_file_pid else "")
            Logger.redirect_all_logging_to_file(name)
        if self.args.log_mongo:
            Logger.duplicate_all_logging_to_mongo(self.args.log_mongo, self.log_id, self.id)
        if not self.is_master and self.device is None:
            from .backends import Device
            self.device = Device(backend=kwargs.pop("backend", None) or getattr(self.args, "backend", None),
                                 device=getattr(self.args, "device", ""))
        if self.is_master:
            from .server import Server
            self.agent = Server(self.args.listen_address, self.workflow, launcher=self)
        elif self.is_slave:
            from .client import Client
            self.agent = Client(self.args.master_address, self.workflow, launcher=self)
        if not self.args.matplotlib_backend == "" and not get(root.common.disable.plotting, False):
            try:
                from .graphics_server import GraphicsServer
                self.graphics = GraphicsServer.launch(self)
            except Exception as e:  # noqa: BLE001 - plotting is optional
                self.debug("plotting is disabled: %s", e)
        self.event("work", "begin", mode=self.mode)
        self.dp_ = self._setup_data_parallel()
        self.workflow.initialize(device=self.device, **kwargs)
        if self.dp_ is not None:
            self.dp_.finalize(self.workflow)
        self._initialized = True
        if self.is_master and getattr(self.args, "yarn_nodes", ""):
            found = self.discover_nodes_from_yarn(self.args.yarn_nodes)
            if found:
                self.args.nodes = ",".join(filter(None, [self.args.nodes] + found))
        if self.is_master and self.args.nodes:
            self.launch_remote_progs()

    def _setup_data_parallel(self):
        """Synchronous data parallelism for a plain ``python -m veles_b200 ...`` run started once per GPU (torchrun or
        ``veles_b200.parallel.supervisor``): ``WORLD_SIZE`` > 1 in the environment turns every rank's workflow into a
        replica whose gradients are exchanged by ``parallel.DataParallel`` ("fused" kernel on GPUs, host all-reduce
        over gloo on the numpy backend).  The configured minibatch is the GLOBAL one: each rank materialises its
        1/world slice, so an N-rank run walks the same minibatch sequence as a single process.  Rank 0 keeps the
        side effects (snapshots, plots, reports)."""
        world = int(os.environ.get("WORLD_SIZE", "1") or 1)
        mode = getattr(self.args, "data_parallel", "auto") or "auto"
        wf = self.workflow
        if world <= 1 or mode == "off" or self.is_master or self.is_slave or not hasattr(wf, "gds"):
            return None
        import torch
        import torch.distributed as dist
        on_gpu = self.device is not None and self.device.exists
        if not dist.is_initialized():
            if on_gpu:
                dist.init_process_group("cpu:gloo,cuda:nccl", device_id=self.device.torch_device)
            else:
                dist.init_process_group("gloo")
        if mode == "auto":
            mode = "fused" if on_gpu else "host"
        from .parallel.dp import DataParallel
        dp = DataParallel(mode=mode)
        dp.auto_finish = True
        loader = getattr(wf, "loader", None)
        if loader is not None:
            loader.shard_rank, loader.shard_world = dp.rank, dp.world_size
        if dp.rank != 0:
            for key in ("snapshotting", "plotting", "publishing"):
                setattr(root.common.disable, key, True)
            if getattr(wf, "result_file", None):
                wf.result_file = None           # every rank holds the same (reduced) results: rank 0 writes them
        dp.attach(wf)
        self.info("data parallel: rank %d of %d, mode %s", dp.rank, dp.world_size, mode)
        return dp

    def run(self):
        """Run the workflow and block until it finishes."""
        if not self._initialized:
            raise RuntimeError("Launcher.run() before initialize()")
        self._start_time = time.time()
        self._running = True
        self._finished.clear()
        from .thread_pool import ThreadPool
        ThreadPool.manhole_context = {"workflow": self.workflow, "launcher": self}
        if getattr(self.args, "manhole", False):
            from . import manhole
            self.info("manhole listening on %s (python -m veles_b200.manhole %d)",
                      manhole.install(workflow=self.workflow, launcher=self), os.getpid())
        if not self.args.stealth and get(root.common.web.host, None) and self.is_master:
            self._start_notifier()
        try:
            if self.agent is not None:
                self.agent.run()          # master: serve jobs; slave: request jobs — until the workflow ends
            else:
                self.workflow._sync = False
                self.thread_pool.callInThread(self.workflow.run)
                self._wait()
        finally:
            self._running = False
            self.event("work", "end")
        if self.args.pdb_on_finish:
            import pdb
            pdb.set_trace()

    def _wait(self):
        while not self._finished.wait(0.2):
            if self.thread_pool.failure is not None:
                raise self.thread_pool.failure
            if self.stopped:
                break
        if self.thread_pool.failure is not None:
            raise self.thread_pool.failure

    def boot(self, **kwargs):
        self.initialize(**kwargs)
        self.run()

    def stop(self):
        self.stopped = True
        if self.workflow is not None and self.workflow.is_running:
            self.workflow.stop()
        if self.agent is not None:
            self.agent.stop()
        for p in self._children:
            if p.poll() is None:
                p.terminate()
        self._finished.set()
        if self._thread_pool is not None:
            self._thread_pool.shutdown()
            self._thread_pool = None
        if self.workflow is not None and self._initialized:
            try:
                self.workflow.print_stats()
            except Exception:  # noqa: BLE001
                pass

    def pause(self):
        self.thread_pool.pause()

    def resume(self):
        self.thread_pool.resume()

    # -- slaves ----------------------------------------------------------------------------------------------------------
    def discover_nodes_from_yarn(self, address, timeout=5.0):
        """Node specs ``host/auto:0`` of the RUNNING nodes a YARN ResourceManager lists
        (reference: ``veles/launcher.py:887-906``: GET /ws/v1/cluster/nodes, one slave per ``nodeHostName``).  An
        unreachable manager or an unexpected answer is a warning, not an error: the run goes on with ``--nodes``."""
        import urllib.request
        if ":" not in address.split("//")[-1]:
            address += ":8088"
        if not address.startswith(("http://", "https://")):
            address = "http://" + address
        url = address.rstrip("/") + "/ws/v1/cluster/nodes"
        self.debug("Requesting GET %s", url)
        try:
            with urllib.request.urlopen(url, timeout=timeout) as resp:
                tree = json.loads(resp.read().decode())
            nodes = (tree.get("nodes") or {}).get("node") or []
            hosts = []
            for node in nodes:
                if node.get("state", "RUNNING") != "RUNNING":
                    continue
                host = node.get("nodeHostName")
                if host and host not in hosts:
                    hosts.append(host)
        except Exception as e:  # noqa: BLE001
            self.warning("Failed to get the nodes list from YARN ResourceManager %s: %s", url, e)
            return []
        self.info("YARN ResourceManager lists %d running node(s): %s", len(hosts), ", ".join(hosts))
        return ["%s/auto:0" % h for h in hosts]

    def launch_remote_progs(self):
        """Start slaves described by ``--nodes host/backend:devs[xN],...`` (local ones via subprocess,
        one per device ordinal with CUDA_VISIBLE_DEVICES; remote ones via ssh)."""
        from .backends import Device
        host, port = self.args.listen_address.rsplit(":", 1)
        master = "%s:%s" % (host if host not in ("", "0.0.0.0") else "127.0.0.1", port)
        base = filter_argv(sys.argv[1:], "-l=", "--listen-address=", "-n=", "--nodes=", "-b", "--background",
                           "-m=", "--master-address=", "-d=", "--device=", "-a=", "--backend=")
        launched = 0
        for node in filter(None, self.args.nodes.split(",")):
            nhost, _, spec = node.partition("/")
            backend, _, devs = (spec or "auto").partition(":")
            for dev in (Device.iterparse(devs) if devs else [""]):
                if self.args.max_nodes and launched >= self.args.max_nodes:
                    return
                cmd = [sys.executable, "-m", "veles_b200", "-m", master, "-a", backend] + \
                    (["-d", "0"] if dev != "" else []) + base
                env = dict(os.environ)
                if dev != "":
                    env["CUDA_VISIBLE_DEVICES"] = dev
                if nhost in ("localhost", "127.0.0.1", socket.gethostname(), ""):
                    if self.args.slave_launch_transform:
                        cmd = (self.args.slave_launch_transform % " ".join(cmd)).split()
                    self.info("Launching local slave: %s", " ".join(cmd))
                    self._children.append(subprocess.Popen(cmd, env=env))
                else:
                    remote = "cd %s && CUDA_VISIBLE_DEVICES=%s %s" % (os.getcwd(), dev, " ".join(cmd))
                    self.info("Launching remote slave on %s: %s", nhost, remote)
                    self._children.append(subprocess.Popen(["ssh", nhost, remote]))
                launched += 1

    # -- status heartbeat --------------------------------------------------------------------------------------------------
    def _start_notifier(self):
        def loop():
            import urllib.request
            url = "http://%s:%d/update" % (get(root.common.web.host, "127.0.0.1"), get(root.common.web.port, 8090))
            interval = get(root.common.web.notification_interval, 1)
            while self._running and not self.stopped:
                try:
                    req = urllib.request.Request(url, data=json.dumps(self.status()).encode(),
                                                 headers={"Content-Type": "application/json"})
                    urllib.request.urlopen(req, timeout=1).read()
                except Exception:  # noqa: BLE001 - the status server is optional
                    pass
                time.sleep(interval)
        self._notify_thread = threading.Thread(target=loop, name="status-notifier", daemon=True)
        self._notify_thread.start()

    def _graph_svg(self, wf):
        if wf is None or not hasattr(wf, "generate_graph_svg"):
            return ""
        try:
            return wf.generate_graph_svg()
        except Exception as e:  # noqa: BLE001 - a picture must not break the heartbeat
            self.debug("workflow graph picture failed: %s", e)
            return ""

    def status(self):
        wf = self.workflow
        from .logger import events
        return {"id": self.id, "log_id": self.log_id, "session": events.session, "name": wf.name if wf else None,
                "master": socket.gethostname(),
                "time": time.time() - (self._start_time or time.time()), "user": os.environ.get("USER", "?"),
                "mode": self.mode, "graph": wf.generate_graph(write_on_disk=False) if wf else "",
                "graph_svg": self._graph_svg(wf),
                "slaves": self.agent.slaves_status() if self.is_master and self.agent is not None else {},
                "plots": self.plots_endpoints, "description": (wf.__doc__ or "") if wf else ""}
