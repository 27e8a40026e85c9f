"""Reader — the drop-in for python/memb/reader.py:5-111 of the reference.

Every lookup runs on the GPU through `_memb` (C ABI -> sm_100a kernels); there is no
CPU decode.  Same names, arguments and exceptions as the reference; additive API (not
in the reference): `device=` constructor keyword, `resolve`, `batch_embedding_ids`,
`decode_range`, `to_tensor`, `load_stats`; the keyword `as_tensor=True` on every lookup
returns a `torch.Tensor` that stays on the reader's GPU (the consumer is
an embedding layer on that GPU), and `words` may also be a NumPy 'S'/'U' array or a
`(blob, offsets)` pair of UTF-8 bytes and uint64 offsets (no per-word
Python objects).
"""
from abc import ABC, abstractmethod

from . import _memb


def _cuda(device):
    import torch
    return torch.device('cuda', device)


def _device_rows(device, rows, dim):
    import torch
    return torch.empty((rows, dim), dtype=torch.float32, device=_cuda(device))


def _current_stream(device):
    import torch
    return torch.cuda.current_stream(_cuda(device)).cuda_stream


def _gensim_keyed_vectors():
    try:
        from gensim.models import KeyedVectors
    except ImportError:
        raise ImportError('You must install gensim for KeyedVectors export')
    return KeyedVectors


def tokenizer_word_list(tokenizer):
    '''Words of a keras Tokenizer as a dense list indexed by token id: what
    python/memb/reader.py:100-109 hands to batch_embedding.  Ids run from 1, so slot 0 --
    like every id the tokenizer does not use -- holds '' (looked up like any word: a zero
    row unless '' is in the vocabulary).  With `num_words` set only ids below it are kept
    and the list has exactly `num_words` slots.'''
    vocabulary = tokenizer.word_index
    limit = tokenizer.num_words
    if limit is None:
        size = max(vocabulary.values()) + 1
        kept = vocabulary.items()
    else:
        size = limit
        kept = [(word, token) for word, token in vocabulary.items() if token < limit]
    slots = [''] * size
    for word, token in kept:
        slots[token] = word
    return slots


class BaseReader(ABC):
    '''What Reader and ReadersUnion have in common (python/memb/reader.py:5-46).'''

    def __getitem__(self, key):
        '''reader[word] -> 1-D vector, reader[[words]] -> 2-D matrix.'''
        for kind, lookup in ((str, self.word_embedding), (list, self.batch_embedding)):
            if isinstance(key, kind):
                return lookup(key)
        raise TypeError('Key type is not supported')

    def to_keyed_vectors(self):
        '''The whole model as a gensim KeyedVectors object (python/memb/reader.py:19-30).'''
        keyed_vectors = _gensim_keyed_vectors()(self.dim)
        vocabulary = self.keys()
        keyed_vectors.add(vocabulary, self._all_rows(vocabulary))
        return keyed_vectors

    def _all_rows(self, vocabulary):
        # one row per entry of keys(), in that order
        return self.batch_embedding(vocabulary)

    @abstractmethod
    def keys(self):
        pass

    @abstractmethod
    def word_embedding(self, word):
        pass

    @abstractmethod
    def batch_embedding(self, words):
        pass

    @abstractmethod
    def tokenizer_embedding(self, tokenizer):
        pass


class Reader(BaseReader):
    '''Obtains embeddings for requested words, decoding them on the GPU.

    Parameters
    ----------
    filename : str or pathlib.Path
    num_threads : int
        Host threads for the one-time parse of the file (0 = all cores).  The
        reference uses it for CPU decoding; there is no CPU decoding here.
    device : int, keyword only, additive
        CUDA ordinal to load the tables onto.

    Attributes
    ----------
    dim : int
        Embeddings dimension
    '''

    def __init__(self, filename, num_threads=0, *, device=0):
        super().__init__()
        self._impl = _memb.Reader(str(filename), num_threads, device)

    @property
    def dim(self):
        return self._impl.dim()

    def keys(self):
        '''List of words contained in the model (row-id order).'''
        return self._impl.keys()

    def word_embedding(self, word, as_tensor=False):
        '''1-D float32 array for a word; zeros when the word is absent.'''
        if as_tensor:
            return self.batch_embedding([word], as_tensor=True)[0]
        return self._impl.word_embedding(word)

    def batch_embedding(self, words, as_tensor=False):
        '''2-D float32 array for a list of words; absent words give zero rows.'''
        if not as_tensor:
            return self._impl.batch_embedding(words)
        # marshal once; resolve + decode are queued on torch's CURRENT stream, so they are
        # ordered after whatever still uses the block the caching allocator hands out
        packed = words if isinstance(words, tuple) else _memb.pack_words(words)
        device = self._impl.device()
        out = _device_rows(device, len(packed[1]) - 1, self.dim)
        self._impl.batch_embedding_into(packed, out.data_ptr(), out.shape[0], True, _current_stream(device))
        return out

    def tokenizer_embedding(self, tokenizer, as_tensor=False):
        '''Weights of an Embedding layer for a keras Tokenizer.'''
        return self.batch_embedding(tokenizer_word_list(tokenizer), as_tensor=as_tensor)

    def _all_rows(self, vocabulary):
        # keys() is the row-id order: the full-vocabulary decode needs no word resolution
        return self._impl.decode_range(0, self._impl.rows())

    # ---- additive ----
    @property
    def rows(self):
        return self._impl.rows()

    def resolve(self, words):
        '''int32 row ids (index into keys()), -1 for absent words.'''
        return self._impl.resolve(words)

    def batch_embedding_ids(self, ids, as_tensor=False):
        '''Rows for int32 row ids; negative / out-of-range ids give zero rows.
        With as_tensor=True `ids` may be a CUDA int32 tensor: nothing crosses PCIe and
        the decode is queued on torch's current stream.'''
        if not as_tensor:
            return self._impl.batch_embedding_ids(ids)
        import torch
        device = _cuda(self._impl.device())
        ids = torch.as_tensor(ids, dtype=torch.int32).to(device).contiguous().reshape(-1)
        out = _device_rows(self._impl.device(), ids.numel(), self.dim)
        self._impl.decode_ids_into(ids.data_ptr(), True, ids.numel(), out.data_ptr(), True,
                                   _current_stream(self._impl.device()))
        ids.record_stream(torch.cuda.current_stream(device))
        return out

    def decode_range(self, begin, end, as_tensor=False):
        '''Rows [begin, end) in row-id order as a (end-begin, dim) float32 array.'''
        if not as_tensor:
            return self._impl.decode_range(begin, end)
        if not 0 <= begin <= end <= self._impl.rows():
            raise ValueError('decode_range: rows out of range')
        out = _device_rows(self._impl.device(), end - begin, self.dim)
        self._impl.decode_range_into(begin, end, out.data_ptr(), True, _current_stream(self._impl.device()))
        return out

    def to_tensor(self):
        '''The whole table, row-id order (= keys() order), as a CUDA tensor.'''
        return self.decode_range(0, self._impl.rows(), as_tensor=True)

    def load_stats(self):
        return self._impl.load_stats()
