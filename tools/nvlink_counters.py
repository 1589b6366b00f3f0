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
            # which form the driver answers differs between releases: all links at once (scope UINT_MAX) or link by link,
            # payload ("DATA") or raw counters
            tried = []
            for kind, (tx, rx) in (("data", (pynvml.NVML_FI_DEV_NVLINK_THROUGHPUT_DATA_TX, pynvml.NVML_FI_DEV_NVLINK_THROUGHPUT_DATA_RX)),
                                   ("raw", (pynvml.NVML_FI_DEV_NVLINK_THROUGHPUT_RAW_TX, pynvml.NVML_FI_DEV_NVLINK_THROUGHPUT_RAW_RX))):
                for scope_name, scopes in (("all", [0xFFFFFFFF]), ("per-link", list(range(18)))):
                    self.fields = [(tx, sc) for sc in scopes] + [(rx, sc) for sc in scopes]
                    self.nscopes = len(scopes)
                    try:
                        self.read()
                        self.ok, self.kind = True, "%s/%s" % (kind, scope_name)
                        break
                    except Exception as e:
                        tried.append("%s/%s: %s" % (kind, scope_name, e))
                if self.ok:
                    break
            if not self.ok:
                self.err = "; ".join(tried)
        except Exception as e:          # containers may hide NVML or the counters
            self.err = repr(e)

    def read(self):
        vals = self.nv.nvmlDeviceGetFieldValues(self.h, self.fields)
        tot, good, last = [0, 0], 0, 0
        for i, v in enumerate(vals):
            if v.nvmlReturn != 0:                         # per-link form: links that do not exist answer with an error
                last = v.nvmlReturn
                continue
            good += 1
            tot[i // self.nscopes] += int(v.value.ullVal) * 1024      # KiB -> bytes
        if good == 0:
            raise RuntimeError("NVML return code %d" % last)
        return tot                                       # [tx_bytes, rx_bytes] since the driver started counting

    def delta(self, before):
        now = self.read()
        return [now[0] - before[0], now[1] - before[1]]
