"""Reader — mirrors python/memb/reader.py:5-111 of the reference.

Every lookup runs on the GPU through `_memb` (C ABI -> sm_100a kernels).
Additive API (not in the reference): `device=` constructor keyword,
`resolve`, `batch_embedding_ids`, `decode_range`, `load_stats`; the keyword
`as_tensor=True` on every lookup returns a `torch.Tensor` that stays on the
reader's GPU (the consumer is an embedding layer on that GPU),
and `words` may also be a NumPy 'S'/'U' array or a `(blob, offsets)` pair of
UTF-8 bytes and uint64 offsets (no per-word Python objects).
"""
from abc import ABC, abstractmethod

from . import _memb


def _device_rows(device, rows, dim):
    import torch
    return torch.empty((rows, dim), dtype=torch.float32, device=torch.device('cuda', device))


class BaseReader(ABC):
    def __getitem__(self, key):
        '''Vector for a word (1-D) or a list of words (2-D).'''
        if isinstance(key, str):
            return self.word_embedding(key)
        elif isinstance(key, list):
            return self.batch_embedding(key)
        else:
            raise TypeError('Key type is not supported')

    def to_keyed_vectors(self):
        '''Export model content to a gensim KeyedVectors object.'''
        try:
            from gensim.models import KeyedVectors
        except ImportError:
            raise ImportError('You must install gensim for KeyedVectors export')

        keyed_vectors = KeyedVectors(self.dim)
        words = self.keys()
        keyed_vectors.add(words, self._all_rows(words))
        return keyed_vectors

    def _all_rows(self, words):
        return self.batch_embedding(words)

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


def tokenizer_word_list(tokenizer):
    '''Dense word list indexed by keras token id (python/memb/reader.py:100-109):
    index 0 and every gap hold '' (which is looked up like any other word).'''
    word_indices = tokenizer.word_index.items()
    if tokenizer.num_words is not None:
        word_indices = [item for item in word_indices if item[1] < tokenizer.num_words]
        max_index = tokenizer.num_words
    else:
        max_index = max([item[1] for item in word_indices]) + 1

    sorted_word_list = [''] * max_index
    for word, idx in word_indices:
        sorted_word_list[idx] = word
    return sorted_word_list


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
        out = _device_rows(self._impl.device(), _memb.Reader.count_words(words), self.dim)
        self._impl.batch_embedding_into(words, out.data_ptr(), out.shape[0], True)
        return out

    def tokenizer_embedding(self, tokenizer, as_tensor=False):
        '''Weights of an Embedding layer for a keras Tokenizer.'''
        return self.batch_embedding(tokenizer_word_list(tokenizer), as_tensor=as_tensor)

    def _all_rows(self, words):
        # full-vocabulary decode needs no word resolution: rows 0..V-1
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
        device = torch.device('cuda', self._impl.device())
        ids = torch.as_tensor(ids, dtype=torch.int32).to(device).contiguous().reshape(-1)
        out = _device_rows(self._impl.device(), ids.numel(), self.dim)
        with torch.cuda.device(device):
            stream = torch.cuda.current_stream().cuda_stream
            self._impl.decode_ids_into(ids.data_ptr(), True, ids.numel(), out.data_ptr(), True, stream)
        ids.record_stream(torch.cuda.current_stream(device))
        return out

    def decode_range(self, begin, end, as_tensor=False):
        '''Rows [begin, end) in row-id order as a (end-begin, dim) float32 array.'''
        if not as_tensor:
            return self._impl.decode_range(begin, end)
        import torch
        if not 0 <= begin <= end <= self._impl.rows():
            raise ValueError('decode_range: rows out of range')
        out = _device_rows(self._impl.device(), end - begin, self.dim)
        with torch.cuda.device(out.device):
            self._impl.decode_range_into(begin, end, out.data_ptr(), True, torch.cuda.current_stream().cuda_stream)
        return out

    def to_tensor(self):
        '''The whole table, row-id order (= keys() order), as a CUDA tensor.'''
        return self.decode_range(0, self._impl.rows(), as_tensor=True)

    def load_stats(self):
        return self._impl.load_stats()
