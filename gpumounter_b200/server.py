"""ref: pkg/server/gpu-mount/server.go + the cluster-facing half of pkg/util/gpu/allocator/allocator.go.

The worker's two RPC handlers as a host-side mirror over the C ABI: same decisions, same result enums, same error
strings; the node-local work goes through the native library (batched mount, in-pod probe).  Everything that talks to
Kubernetes or the kubelet is behind the small `Cluster` interface — the reference hard-wires client-go and the
PodResources socket (ref: pkg/config/config.go:30-45, collector.go:165-194), which is control plane and out of this
repository's scope; tests drive the handlers with a fake cluster, and `serve()` puts them behind the reference's exact
gRPC method paths.

Differences from the reference, all deliberate and none visible on the wire:
  * the slave-pod polls sleep between rounds instead of spinning on the API server (ref: allocator.go:246-281,295-316);
  * `gpu_num == 0` with `is_entire_mount` is refused instead of dividing by zero (ref: server.go:61-65, allocator.go:49);
  * an optional post-mount probe runs before Success (hook point ref: server.go:97-98).
"""

from __future__ import annotations

import threading
import time
from concurrent import futures
from dataclasses import dataclass
from typing import Callable, List, Optional, Sequence

from . import allocator, api, device, util
from ._lib import GMP_ERR_BUSY, ProbeError
from ._model import MountEnv, Pod
from .collector import GPUPoolNamespace, Listing

# ref: pkg/util/gpu/types.go:12-16
InsufficientGPU = "InsufficientGPU"
SuccessfullyCreated = "SuccessfullyCreated"
FailedCreated = "FailedCreated"
SuccessfullyDeleted = "SuccessfullyDeleted"
FailedDeleted = "FailedDeleted"
ServiceInternalError = "Service Internal Error "  # trailing space as in the reference (server.go:41)
PodReasonUnschedulable = "Unschedulable"


class NotFound(Exception):
    """k8s_errors.IsNotFound"""


class ClusterError(Exception):
    """any other API-server / kubelet failure"""


class RpcError(Exception):
    """What the Go handler returns as `nil, err`: gRPC status Unknown with this message."""


@dataclass
class PodStatus:
    phase: str                                   # pod.Status.Phase
    first_condition_reason: Optional[str] = None  # pod.Status.Conditions[0].Reason, None when there are no conditions


class Cluster:
    """What the handlers need from Kubernetes and the kubelet."""

    def get_pod(self, namespace: str, name: str) -> Pod:  # raises NotFound / ClusterError
        raise NotImplementedError

    def create_slave_pod(self, name: str, owner: Pod, gpu_num: int) -> None:
        """ref: allocator.go:189-234 — ns gpu-pool, label app=gpu-pool, image alpine:latest, limits nvidia.com/gpu=gpu_num,
        nodeSelector kubernetes.io/hostname=<owner node>, ownerReference to the owner pod."""
        raise NotImplementedError

    def pod_status(self, namespace: str, name: str) -> PodStatus:  # raises NotFound / ClusterError
        raise NotImplementedError

    def delete_pod(self, namespace: str, name: str, grace_period_seconds: Optional[int] = None) -> None:
        raise NotImplementedError

    def list_pod_resources(self) -> Listing:
        """kubelet v1alpha1 PodResourcesLister/List (ref: collector.go:182-194)"""
        raise NotImplementedError

    def wait_for_pod_event(self, namespace: str, names: Sequence[str], timeout_s: float) -> bool:
        """Extension (the reference spins on Get, allocator.go:246,295): block until one of the pods
        changes - a watch on the namespace - or `timeout_s` passes, and return True.  The default says "no watch here"
        (False) and the caller sleeps between rounds instead."""
        return False


ProbeHook = Callable[..., "tuple[int, dict]"]  # (pod, newly mounted GPUs[, GPUs the pod held before]) -> (status, report)


class GPUMountImpl:
    """ref: server.go:17-32"""

    def __init__(self, cluster: Cluster, gpus: Optional[Sequence[device.NvidiaGPU]] = None, env: Optional[MountEnv] = None,
                 probe_hook: Optional[ProbeHook] = None, poll_interval_s: float = 0.002, poll_timeout_s: Optional[float] = None,
                 extended_response: bool = False, fail_on_probe_error: bool = False, revoke_on_mount_failure: bool = True):
        self.cluster = cluster
        # additive wire surface: answer AddGPU with the probe verdict in NEW field numbers (2..4);
        # a peer built from the reference's api.proto skips them as unknown fields and sees the same enum in field 1
        self.extended_response = extended_response
        # Policy for the post-mount probe (no reference counterpart).  False: a failing probe is reported (probe_status in the
        # extended response, last_probe_report) and AddGPU still answers Success, as the reference would have.  True: the
        # GPUs are unmounted again, the slave pods deleted and the RPC fails with ServiceInternalError.
        self.fail_on_probe_error = fail_on_probe_error
        # The batched mount grants ALL n devices.allow lines before the first mknod; if it fails part-way the reference's
        # rollback (delete the slave pods, server.go:86-91) would hand GPUs back to the scheduler that the pod can still
        # open.  True: permissions and nodes are revoked before the slave pods go.  False: exactly the reference's rollback.
        self.revoke_on_mount_failure = revoke_on_mount_failure
        # gRPC serves RPCs concurrently and UpdateGPUStatus rewrites the shared GPU list in place (the reference has the same
        # race, collector.go:113-135,140-144): every refresh-then-query pair runs under this lock
        self._status_lock = threading.Lock()
        self.env = env
        self.probe_hook = probe_hook
        self.poll_interval_s = poll_interval_s
        self.poll_timeout_s = poll_timeout_s
        self.allocator = allocator.GPUAllocator()  # ref: NewGPUAllocator -> NewGPUCollector (allocator.go:27-38)
        if gpus is None:
            self.allocator.GetGPUInfo()             # NVML enumeration, once (collector.go:40-79)
        else:
            self.allocator.GPUList = list(gpus)
        self.allocator.UpdateGPUStatus(cluster.list_pod_resources())  # collector.go:31
        self.last_probe_report: Optional[dict] = None

    # ---- collector / allocator calls that need the kubelet ---------------------------------------------------
    def GetPodGPUResources(self, podName: str, namespace: str) -> List[device.NvidiaGPU]:  # ref: collector.go:149-163
        with self._status_lock:
            self.allocator.UpdateGPUStatus(self.cluster.list_pod_resources())
            return self.allocator.GetPodGPUResources(podName, namespace)

    def GetMountType(self, pod: Pod) -> str:  # ref: allocator.go:158-187
        with self._status_lock:
            try:
                self.allocator.UpdateGPUStatus(self.cluster.list_pod_resources())
            except (ClusterError, ProbeError):
                return util.UnknownMount  # allocator.go:161-165
            return self.allocator.GetMountType(pod.name, pod.namespace)

    def GetRemoveGPU(self, ownerPod: Pod, uuids: Sequence[str]) -> List[device.NvidiaGPU]:  # ref: allocator.go:101-126
        with self._status_lock:
            self.allocator.UpdateGPUStatus(self.cluster.list_pod_resources())
            return self.allocator.GetRemoveGPU(ownerPod.name, ownerPod.namespace, list(uuids))

    # ---- slave pods (ref: allocator.go:40-99, 128-156, 236-317) ------------------------------------------------
    def _poll(self, step: Callable[[], Optional[str]], podNames: Sequence[str] = ()) -> str:
        deadline = None if self.poll_timeout_s is None else time.monotonic() + self.poll_timeout_s
        delay = self.poll_interval_s
        while True:
            verdict = step()
            if verdict is not None:
                return verdict
            if deadline is not None and time.monotonic() > deadline:
                return FailedCreated
            # a cluster with a watch wakes us on the next pod event (re-checked at least once a second);
            # without one: sleep with back-off, where the reference spins
            if not self.cluster.wait_for_pod_event(GPUPoolNamespace, podNames, 1.0):
                time.sleep(delay)
                delay = min(delay * 2, 0.25)

    def checkCreateState(self, podNames: Sequence[str]) -> str:  # ref: allocator.go:236-282
        def one_round() -> Optional[str]:
            flag = True
            for name in podNames:
                try:
                    st = self.cluster.pod_status(GPUPoolNamespace, name)
                except NotFound:
                    continue  # the reference logs "Not Found...." and does NOT clear the flag (allocator.go:251-253)
                except ClusterError:
                    return FailedCreated
                if st.phase == "Running":
                    continue
                if st.first_condition_reason is None:
                    flag = False  # creating
                elif st.first_condition_reason == PodReasonUnschedulable:
                    return InsufficientGPU
                else:
                    flag = False
            return SuccessfullyCreated if flag else None

        return self._poll(one_round, podNames)

    def checkDeleteState(self, podNames: Sequence[str]) -> str:  # ref: allocator.go:284-317
        def one_round() -> Optional[str]:
            flag = True
            for name in podNames:
                try:
                    self.cluster.pod_status(GPUPoolNamespace, name)
                except NotFound:
                    continue
                except ClusterError:
                    return FailedDeleted
                flag = False
            return SuccessfullyDeleted if flag else None

        v = self._poll(one_round, podNames)
        return FailedDeleted if v == FailedCreated else v

    def GetAvailableGPU(self, ownerPod: Pod, totalGpuNum: int, gpuNumPerPod: int) -> List[device.NvidiaGPU]:
        # ref: allocator.go:40-99; raises RpcError(InsufficientGPU | FailedCreated)
        if gpuNumPerPod == 0:
            raise RpcError(FailedCreated)  # Go: integer divide by zero panic at allocator.go:49
        names: List[str] = []
        for _ in range(max(0, int(totalGpuNum / gpuNumPerPod))):  # Go's integer division truncates toward zero
            name = allocator.newGPUSlavePodName(ownerPod.name)
            try:
                self.cluster.create_slave_pod(name, ownerPod, gpuNumPerPod)
            except ClusterError:
                raise RpcError(FailedCreated)  # allocator.go:53-57: already created pods are NOT cleaned up here
            names.append(name)
        verdict = self.checkCreateState(names)
        if verdict in (InsufficientGPU, FailedCreated):  # allocator.go:65-83
            for name in names:
                try:
                    self.cluster.delete_pod(GPUPoolNamespace, name, grace_period_seconds=0)
                except (ClusterError, NotFound):
                    pass  # logged and ignored in the reference
            raise RpcError(verdict)
        out: List[device.NvidiaGPU] = []
        for name in names:  # allocator.go:86-94
            try:
                out.extend(self.GetPodGPUResources(name, GPUPoolNamespace))
            except (ClusterError, ProbeError):
                raise RpcError(FailedCreated)
        return out

    def DeleteSlavePods(self, slavePodNames: Sequence[str]) -> None:  # ref: allocator.go:128-156
        for name in slavePodNames:
            try:
                self.cluster.delete_pod(GPUPoolNamespace, name, grace_period_seconds=None)  # default DeleteOptions{}
            except ClusterError as e:
                raise RpcError(str(e))
        if self.checkDeleteState(slavePodNames) != SuccessfullyDeleted:
            raise RpcError("Failed to delete slave pods ")

    # ---- the two RPCs --------------------------------------------------------------------------------------------
    def AddGPU(self, request) -> "api.AddGPUResponse":  # ref: server.go:34-99
        try:
            targetPod = self.cluster.get_pod(request.namespace, request.pod_name)
        except NotFound:
            return api.AddGPUResponse(add_gpu_result=api.AddGPUResult_PodNotFound)  # server.go:45-48
        except ClusterError:
            raise RpcError(ServiceInternalError)
        if not util.CanMount(self.GetMountType(targetPod), request.is_entire_mount):
            raise RpcError(FailedCreated)  # server.go:57-59
        gpuNum = int(request.gpu_num)
        gpuNumPerPod = gpuNum if request.is_entire_mount else 1  # server.go:61-65
        try:
            gpuResources = self.GetAvailableGPU(targetPod, gpuNum, gpuNumPerPod)
        except RpcError as e:
            if str(e) == InsufficientGPU:
                return api.AddGPUResponse(add_gpu_result=api.AddGPUResult_InsufficientGPU)  # server.go:69-71
            raise RpcError(ServiceInternalError)  # server.go:72-77
        def free_slave_pods():  # server.go:86-91
            for freeGPU in gpuResources:
                try:
                    self.cluster.delete_pod(GPUPoolNamespace, freeGPU.PodName, grace_period_seconds=0)
                except (ClusterError, NotFound):
                    pass

        def revoke_all():
            for g in gpuResources:  # one by one: some have a node, all have a devices.allow line, any step may fail
                try:
                    util.UnmountGPU(targetPod, g, True, self.env)
                except ProbeError:
                    pass

        try:
            util.MountGPUs(targetPod, gpuResources, self.env)  # the loop of server.go:80-95, batched
        except ProbeError:
            if self.revoke_on_mount_failure:
                revoke_all()
            free_slave_pods()
            raise RpcError(ServiceInternalError)
        status, report = None, None
        if self.probe_hook is not None:  # new: prove the mounted GPUs usable from inside the pod
            try:
                try:  # what the pod held before this call (a hook that takes a third argument is told)
                    new = {g.UUID for g in gpuResources}
                    already = [g for g in self.GetPodGPUResources(targetPod.name, targetPod.namespace) if g.UUID not in new]
                except (ClusterError, ProbeError):
                    already = []
                import inspect

                if len(inspect.signature(self.probe_hook).parameters) >= 3:
                    status, report = self.probe_hook(targetPod, gpuResources, already)
                else:
                    status, report = self.probe_hook(targetPod, gpuResources)
            except ProbeError as e:  # a hook that raises (GMP_ERR_TIMEOUT included) is a failed probe, not a crashed RPC
                status, report = e.code, {"detail": e.detail}
            except Exception as e:  # noqa: BLE001 - whatever the hook does, AddGPU keeps its contract
                status, report = -9, {"detail": f"{type(e).__name__}: {e}"}
            self.last_probe_report = dict(report or {}, status=status)
            if status not in (None, 0) and self.fail_on_probe_error:
                revoke_all()
                free_slave_pods()
                raise RpcError(ServiceInternalError)
        if self.extended_response:
            import json as _json

            return api.ExtendedAddGPUResponse(add_gpu_result=api.AddGPUResult_Success,
                                              probe_status=0 if status is None else int(status),
                                              probe_report="" if report is None else _json.dumps(report, separators=(",", ":")),
                                              mounted_uuids=[g.UUID for g in gpuResources])
        return api.AddGPUResponse(add_gpu_result=api.AddGPUResult_Success)

    def RemoveGPU(self, request) -> "api.RemoveGPUResponse":  # ref: server.go:101-179
        try:
            targetPod = self.cluster.get_pod(request.namespace, request.pod_name)
        except NotFound:
            return api.RemoveGPUResponse(remove_gpu_result=api.RemoveGPUResult_PodNotFound)
        except ClusterError:
            raise RpcError(ServiceInternalError)
        try:
            removeGPUs = self.GetRemoveGPU(targetPod, list(request.uuids))
        except (ClusterError, ProbeError) as e:
            raise RpcError(str(e))
        if len(removeGPUs) == 0:
            return api.RemoveGPUResponse(remove_gpu_result=api.RemoveGPUResult_GPUNotFound)  # server.go:130-135
        slavePodNames = []
        # the busy check asks NVML once per GPU here and once more inside the unmount (each init .. shutdown, as
        # nvidia.go:58-63): one attach held across the handler makes those calls cheap (30-140 ms each otherwise)
        with util.NvmlAttached():
            for g in removeGPUs:  # server.go:139-153
                slavePodNames.append(g.PodName)
                try:
                    procs = util.GetPodGPUProcesses(targetPod, g, self.env)
                except ProbeError as e:
                    raise RpcError(str(e))
                if procs is not None and not request.force:
                    return api.RemoveGPUResponse(remove_gpu_result=api.RemoveGPUResult_GPUBusy)
            try:
                util.UnmountGPUs(targetPod, removeGPUs, bool(request.force), self.env)  # server.go:155-168, batched
            except ProbeError as e:
                if e.code == GMP_ERR_BUSY:  # the GPUBusy sentinel of util.go:108 / server.go:158
                    return api.RemoveGPUResponse(remove_gpu_result=api.RemoveGPUResult_GPUBusy)
                raise RpcError(str(e))
        self.DeleteSlavePods(slavePodNames)  # server.go:171-175
        return api.RemoveGPUResponse(remove_gpu_result=api.RemoveGPUResult_Success)


def in_pod_probe_hook(env: Optional[MountEnv] = None, probe_args: Sequence[str] = ("--bytes", "1G", "--iters", "1"),
                      nodes_first: bool = True) -> "ProbeHook":
    """The product's post-mount hook (insertion point ref: server.go:97-98), for
    GPUMountImpl(probe_hook=...).  Stage 1 (nothing of the pod executed): the device nodes open from inside the pod's
    device cgroup and mount namespace.  Stage 2: gpumounter-probe inside the pod - one process per GPU from two GPUs on
    (gmp_probe_in_pod_fanout, cgroup v1), one process otherwise.  Returns (status, report); never raises."""
    from ._model import EXT_CGROUP_V2

    def hook(pod: Pod, gpus: Sequence[device.NvidiaGPU], already: Sequence[device.NvidiaGPU] = ()):
        # `already`: GPUs the pod held before this AddGPU.  CUDA in the pod must see them too (the visible set has to be
        # exactly new + old), but they are not probed - the tenant may be using them.
        args = list(probe_args)
        for g in already:
            args += ["--also-mounted", f"{g.UUID}:{g.MinorNumber}"]
        if nodes_first:
            rc, failed = util.CheckNodesInPod(pod, gpus, env, raise_on_error=False)
            if rc != 0:
                from ._lib import lib

                return rc, {"stage": "nodes", "failed_index": failed, "detail": lib().gmp_last_error().decode(errors="replace")}
        v2 = env is not None and bool(env.extensions & EXT_CGROUP_V2)
        fn = util.ProbeInPodFanout if (len(gpus) >= 2 and not v2) else util.ProbeInPod
        return fn(pod, gpus, args, env, raise_on_error=False)

    return hook


def serve(impl: GPUMountImpl, address: str = f"[::]:{api.WORKER_PORT}", max_workers: int = 8):
    """Both services on one insecure port, as ref: cmd/GPUMounter-worker/main.go:24-33.  Returns the started grpc.Server."""
    import grpc

    def wrap(fn, req_cls):
        def handler(raw_request, context):
            try:
                return fn(raw_request)
            except RpcError as e:  # Go: `return nil, err` -> status Unknown, message = err.Error()
                context.abort(grpc.StatusCode.UNKNOWN, str(e))

        return grpc.unary_unary_rpc_method_handler(handler, request_deserializer=req_cls.FromString,
                                                   response_serializer=lambda m: m.SerializeToString())

    server = grpc.server(futures.ThreadPoolExecutor(max_workers=max_workers))
    server.add_generic_rpc_handlers((
        grpc.method_handlers_generic_handler("gpu_mount.AddGPUService", {"AddGPU": wrap(impl.AddGPU, api.AddGPURequest)}),
        grpc.method_handlers_generic_handler("gpu_mount.RemoveGPUService",
                                             {"RemoveGPU": wrap(impl.RemoveGPU, api.RemoveGPURequest)}),
    ))
    port = server.add_insecure_port(address)
    server.start()
    server.bound_port = port
    return server
