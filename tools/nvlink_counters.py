"""NVLink byte counters of one GPU through NVML (NVML_FI_DEV_NVLINK_THROUGHPUT_DATA_TX / _RX summed over its links, KiB),
read around a timed region so that an NVLink-bound number comes with the traffic the links actually carried."""
import torch


class NvlinkCounters(object):
    def __init__(self, dev_index):
        self.ok, self.err = False, None
        try:
            import pynvml
            pynvml.nvmlInit()
            pr = torch.cuda.get_device_properties(dev_index)
            bdf = "%08x:%02x:%02x.0" % (pr.pci_domain_id, pr.pci_bus_id, pr.pci_device_id)
            self.nv, self.h = pynvml, pynvml.nvmlDeviceGetHandleByPciBusId(bdf.encode())
            self.fields = [(pynvml.NVML_FI_DEV_NVLINK_THROUGHPUT_DATA_TX, 0xFFFFFFFF), (pynvml.NVML_FI_DEV_NVLINK_THROUGHPUT_DATA_RX, 0xFFFFFFFF)]
            self.read()
            self.ok = True
        except Exception as e:          # containers may hide NVML or the counters
            self.err = repr(e)

    def read(self):
        vals = self.nv.nvmlDeviceGetFieldValues(self.h, self.fields)
        out = []
        for v in vals:
            if v.nvmlReturn != 0:
                raise RuntimeError("NVML field %d: return code %d" % (v.fieldId, v.nvmlReturn))
            out.append(int(v.value.ullVal) * 1024)      # KiB -> bytes
        return out                                       # [tx_bytes, rx_bytes] since the driver started counting

    def delta(self, before):
        now = self.read()
        return [now[0] - before[0], now[1] - before[1]]
