"""Host placement for one-process-per-GPU deployments: run the submitting threads, and allocate the pinned LLR / payload
buffers, on the CPU socket the GPU's PCIe root port hangs off.

With host-resident LLRs the chain is PCIe-bound; on a multi-socket host a rank whose pinned buffers live on the
other socket pays the inter-socket link on every copy. The reference pins its PHY worker threads for the same reason
(``du_low`` executors / ``cpu_affinities`` in the application configuration). Nothing here touches the GPU.
"""
import os


def _parse_cpulist(text):
    cpus = set()
    for part in text.strip().split(","):
        if not part:
            continue
        a, _, b = part.partition("-")
        cpus.update(range(int(a), int(b or a) + 1))
    return cpus


def device_local_cpus(device):
    """CPUs local to CUDA device `device` (a torch ordinal), from sysfs; None when it cannot be told."""
    import torch

    p = torch.cuda.get_device_properties(device)
    try:
        bdf = f"{p.pci_domain_id:04x}:{p.pci_bus_id:02x}:{p.pci_device_id:02x}.0"
    except AttributeError:
        return None
    try:
        with open(f"/sys/bus/pci/devices/{bdf}/local_cpulist") as f:
            cpus = _parse_cpulist(f.read())
    except OSError:
        return None
    return cpus or None


def bind_to_device_socket(device):
    """Restricts the calling process to the CPUs local to the device (threads created afterwards inherit it, and memory
    pinned afterwards is first-touched on that node). Returns a short description, or None when nothing was changed."""
    cpus = device_local_cpus(device)
    if not cpus:
        return None
    allowed = os.sched_getaffinity(0)
    target = cpus & allowed
    if not target or target == allowed:
        return None
    os.sched_setaffinity(0, target)
    return f"{len(target)} CPUs local to cuda:{device}"
