"""ReadersUnion — mirrors python/memb/readers_union.py:5-104 of the reference.

The merge (np.mean(axis=0) / np.concatenate(axis=-1) in the reference) is fused
into the GPU decode: no per-reader (n, dim) temporaries exist.
"""
import numpy as np

from . import _memb
from .reader import BaseReader, Reader, tokenizer_word_list


class AverageUnionMaker:
    name = 'average'

    @staticmethod
    def check(readers):
        dims = [reader.dim for reader in readers]
        if any(dim != dims[0] for dim in dims):
            raise AssertionError('Dimensions of all readers must be equal for average mode')

    @staticmethod
    def dim(reader_dims):
        return reader_dims[0]

    @staticmethod
    def merge(vectors):
        return np.mean(vectors, axis=0)


class ConcatenatedUnionMaker:
    name = 'concatenate'

    @staticmethod
    def check(readers):
        pass

    @staticmethod
    def dim(reader_dims):
        return sum(reader_dims)

    @staticmethod
    def merge(vectors):
        return np.concatenate(vectors, axis=-1)


UNION_MAKERS = {
    'average': AverageUnionMaker(),
    'concatenate': ConcatenatedUnionMaker(),
}


class ReadersUnion(BaseReader):
    '''Makes a list of Readers behave like one, returning the average or the
    concatenation of their embeddings.

    Parameters
    ----------
    readers : list of Reader
    mode : str
        'average' or 'concatenate'
    '''

    def __init__(self, readers, mode):
        super().__init__()

        if len(readers) < 2:
            raise AssertionError('You must pass at least 2 readers to create a union')

        self._union_maker = UNION_MAKERS.get(mode)
        if self._union_maker is None:
            raise KeyError('Mode {} is not supported. Available modes are {}'.format(
                mode, list(UNION_MAKERS.keys())))

        self._union_maker.check(readers)
        self._readers = readers
        # the fused kernel path needs native readers; anything else (e.g. a
        # nested union) is merged with the reference's NumPy expressions
        self._native = all(isinstance(reader, Reader) for reader in readers)

    @property
    def dim(self):
        return self._union_maker.dim([reader.dim for reader in self._readers])

    def keys(self):
        '''Union of keys contained in wrapped models'''
        all_keys = set()
        for reader in self._readers:
            all_keys |= set(reader.keys())

        return sorted(all_keys)

    def word_embedding(self, word, as_tensor=False):
        '''Merged vectors from all readers for a single word'''
        if self._native:
            return self.batch_embedding([word], as_tensor=as_tensor)[0]
        return self._union_maker.merge([reader.word_embedding(word) for reader in self._readers])

    def batch_embedding(self, words, as_tensor=False):
        '''Merged vectors from all readers for a list of words.
        as_tensor=True (additive, native readers only): CUDA tensor on the readers' GPU.'''
        if self._native:
            impls = [reader._impl for reader in self._readers]
            if as_tensor:
                import torch
                device = torch.device('cuda', impls[0].device())
                out = torch.empty((_memb.Reader.count_words(words), self.dim), dtype=torch.float32, device=device)
                _memb.union_batch_embedding_into(impls, self._union_maker.name, words, out.data_ptr(),
                                                 out.shape[0], True)
                return out
            return _memb.union_batch_embedding(impls, self._union_maker.name, words)
        if as_tensor:
            raise TypeError('as_tensor needs native Readers in the union')
        return self._union_maker.merge([reader.batch_embedding(words) for reader in self._readers])

    def tokenizer_embedding(self, tokenizer, as_tensor=False):
        '''Merged results of tokenizer_embedding from all readers'''
        if self._native:
            return self.batch_embedding(tokenizer_word_list(tokenizer), as_tensor=as_tensor)
        return self._union_maker.merge([reader.tokenizer_embedding(tokenizer) for reader in self._readers])
