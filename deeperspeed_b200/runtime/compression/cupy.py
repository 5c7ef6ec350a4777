"""Bit packing helpers with the API of the reference's CupyBackend (deepspeed/runtime/compression/cupy.py:10-24) but
implemented on torch tensors (cupy is not required): packbits / unpackbits in big-endian bit order like numpy."""
import torch


class CupyBackend:
    def __init__(self):
        pass

    def torch2cupy(self, tensor):
        return tensor

    def cupy2torch(self, cupy_tensor):
        return cupy_tensor

    def compress_by_chunk(self, cupy_bool_tensor, num_chunks):
        packed = packbits(cupy_bool_tensor)
        return list(torch.chunk(packed, num_chunks))


_W = None


def packbits(bool_tensor):
    b = bool_tensor.to(torch.uint8).view(-1, 8)
    w = torch.tensor([128, 64, 32, 16, 8, 4, 2, 1], dtype=torch.uint8, device=b.device)
    return (b * w).sum(dim=1).to(torch.uint8)


def unpackbits(byte_tensor):
    w = torch.tensor([128, 64, 32, 16, 8, 4, 2, 1], dtype=torch.uint8, device=byte_tensor.device)
    return ((byte_tensor.view(-1, 1) & w) != 0).view(-1)
