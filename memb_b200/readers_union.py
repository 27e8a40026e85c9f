"""ReadersUnion — the drop-in for python/memb/readers_union.py:5-104 of the reference.

The reference merges per-reader (n, dim) NumPy batches with np.mean(axis=0) or
np.concatenate(axis=-1) (:16-18, :30-32).  Here a union of Readers is ONE native call:
every reader's words are resolved on the GPU and the merge is fused into the decode
(trained readers: one kernel launch, memb_b200/csrc/trained_rows.cuh), so no per-reader
batch ever exists.  A union with a nested memb_b200 union among its members merges the
members' device tensors on the GPU with the same arithmetic -- fp32 adds in member order,
then one fp32 divide by the member count.  The reference accepts ANY object with the reader
interface; such foreign members hand back host arrays of their own, and only those are
merged on the host (the one host-side merge in the package: nothing of memb_b200's own
lookup path runs there).
"""
import numpy as np

from . import _memb
from .reader import BaseReader, Reader, tokenizer_word_list

# mode -> (dimension of the merged vectors, whether every member must have the same dim)
_MODES = {
    'average': (lambda dims: dims[0], True),
    'concatenate': (sum, False),
}


class ReadersUnion(BaseReader):
    '''Makes a list of Readers behave like one, returning the average or the
    concatenation of their embeddings.

    Parameters
    ----------
    readers : list of Reader
    mode : str
        'average' or 'concatenate'

    Attributes
    ----------
    dim : int
        Dimension of vectors after merge
    '''

    def __init__(self, readers, mode):
        super().__init__()
        # same checks, order and messages as python/memb/readers_union.py:56-69
        if len(readers) < 2:
            raise AssertionError('You must pass at least 2 readers to create a union')
        if mode not in _MODES:
            raise KeyError('Mode {} is not supported. Available modes are {}'.format(mode, list(_MODES.keys())))
        merged_dim, equal_dims = _MODES[mode]
        if equal_dims and len({reader.dim for reader in readers}) > 1:
            raise AssertionError('Dimensions of all readers must be equal for average mode')
        self._mode = mode
        self._merged_dim = merged_dim
        self._readers = readers
        self._native = all(isinstance(reader, Reader) for reader in readers)
        self._on_device = all(isinstance(reader, (Reader, ReadersUnion)) for reader in readers)

    @property
    def dim(self):
        return self._merged_dim([reader.dim for reader in self._readers])

    def keys(self):
        '''Union of keys contained in wrapped models'''
        return sorted(set().union(*(reader.keys() for reader in self._readers)))

    def word_embedding(self, word, as_tensor=False):
        '''Merged vectors from all readers for a single word'''
        return self.batch_embedding([word], as_tensor=as_tensor)[0]

    def batch_embedding(self, words, as_tensor=False):
        '''Merged vectors from all readers for a list of words.
        as_tensor=True (additive): CUDA tensor on the readers' GPU.'''
        if not self._on_device:
            if as_tensor:
                raise TypeError('as_tensor needs memb_b200 Readers or ReadersUnions in the union')
            return self._merge_foreign([reader.batch_embedding(words) for reader in self._readers])
        if not self._native:
            merged = self._merge_on_device(words)
            return merged if as_tensor else merged.cpu().numpy()
        impls = [reader._impl for reader in self._readers]
        if not as_tensor:
            return _memb.union_batch_embedding(impls, self._mode, words)
        import torch
        device = torch.device('cuda', impls[0].device())
        packed = words if isinstance(words, tuple) else _memb.pack_words(words)
        out = torch.empty((len(packed[1]) - 1, self.dim), dtype=torch.float32, device=device)
        _memb.union_batch_embedding_into(impls, self._mode, packed, out.data_ptr(), out.shape[0], True,
                                         torch.cuda.current_stream(device).cuda_stream)
        return out

    def tokenizer_embedding(self, tokenizer, as_tensor=False):
        '''Merged results of tokenizer_embedding from all readers'''
        return self.batch_embedding(tokenizer_word_list(tokenizer), as_tensor=as_tensor)

    def _merge_on_device(self, words):
        # members that are unions themselves: each yields a device tensor, merged here with
        # torch ops on the GPU in the reference's order (sequential fp32 adds, one true divide)
        import torch
        parts = [reader.batch_embedding(words, as_tensor=True) for reader in self._readers]
        if self._mode == 'concatenate':
            return torch.cat(parts, dim=-1)
        total = parts[0].clone()
        for part in parts[1:]:
            total += part
        total /= float(len(parts))
        return total

    def _merge_foreign(self, arrays):
        # host arrays produced by reader objects that are not memb_b200's
        # (python/memb/readers_union.py:16-18,30-32)
        return np.concatenate(arrays, axis=-1) if self._mode == 'concatenate' else np.mean(arrays, axis=0)
