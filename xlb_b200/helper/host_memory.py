"""Pinned host memory placed next to the GPU.

``HostStreamedStepper`` moves whole lattices over PCIe every step, so where the pinned pages live matters: on a two-socket
host a buffer on the far socket crosses the inter-socket link on every transfer.  ``pinned_empty`` allocates (and thereby
first-touches) the buffer while the calling thread is confined to the CPUs that are local to the GPU, read from
``/sys/bus/pci/devices/<bus id>/local_cpulist``; the affinity is restored afterwards.  Where the topology cannot be read
(no sysfs entry, a container without it) this is a plain ``torch.empty(..., pin_memory=True)``.
"""

import os

import torch


def gpu_local_cpus(device=None):
    """CPU ids local to ``device`` (default: the current CUDA device), or None when unknown."""
    try:
        props = torch.cuda.get_device_properties(torch.cuda.current_device() if device is None else device)
        bus = f"{props.pci_domain_id:04x}:{props.pci_bus_id:02x}:{props.pci_device_id:02x}.0"
        with open(f"/sys/bus/pci/devices/{bus}/local_cpulist") as fh:
            text = fh.read().strip()
    except (OSError, AttributeError, RuntimeError):
        return None
    cpus = set()
    for part in text.split(","):
        if not part:
            continue
        lo, _, hi = part.partition("-")
        cpus.update(range(int(lo), int(hi or lo) + 1))
    allowed = os.sched_getaffinity(0) if hasattr(os, "sched_getaffinity") else cpus
    cpus &= set(allowed)
    return cpus or None


def pinned_empty(shape, dtype, device=None):
    """``torch.empty(shape, dtype=dtype, pin_memory=True)`` with its pages on the NUMA node of ``device``.
    Returns ``(tensor, note)``; ``note`` says what placement was applied."""
    cpus = gpu_local_cpus(device)
    if cpus is None or not hasattr(os, "sched_setaffinity"):
        return torch.empty(shape, dtype=dtype, pin_memory=True), "default placement (GPU-local CPU list not available)"
    before = os.sched_getaffinity(0)
    try:
        os.sched_setaffinity(0, cpus)
        t = torch.empty(shape, dtype=dtype, pin_memory=True)
        t.view(-1)[:: max(1, 4096 // t.element_size())] = 0  # first touch of every page from a GPU-local CPU
    finally:
        os.sched_setaffinity(0, before)
    return t, f"first-touched from the {len(cpus)} GPU-local CPUs of {len(before)} allowed"
