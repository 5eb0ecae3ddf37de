"""Corpus preparation helpers.

Ref: src/models/base/pytorchavitm/utils/data_preparation.py:11-64 (``prepare_dataset``),
src/models/base/contextualized_topic_models/utils/data_preparation.py:195-328
(``TopicModelDataPreparation``, ``prepare_ctm_dataset``, ``prepare_hold_out_dataset``) and
utils/preprocessing.py:6-60 (``WhiteSpacePreprocessing``).  Bag-of-words matrices stay sparse.
"""
from __future__ import annotations

import string
from typing import List, Optional, Sequence

import numpy as np
import scipy.sparse as sp
from sklearn.feature_extraction.text import CountVectorizer
from sklearn.model_selection import train_test_split
from sklearn.preprocessing import OneHotEncoder

from .datasets import BOWDataset, CTMDataset


def prepare_dataset(corpus: Sequence[str], val_size: float = 0.25, max_features: Optional[int] = None):
    """75/25 split (seed 42), ``CountVectorizer`` fitted on the training part.
    Returns ``(train_dataset, val_dataset, input_size, id2token, docs_train, cv)``."""
    docs_train, docs_val = train_test_split(list(corpus), test_size=val_size, random_state=42)
    cv = CountVectorizer(input="content", lowercase=True, stop_words="english", binary=False,
                         max_features=max_features)
    train_bow = cv.fit_transform(docs_train)
    val_bow = cv.transform(docs_val)
    idx2token = cv.get_feature_names_out()
    id2token = dict(enumerate(idx2token))
    return (BOWDataset(train_bow, idx2token), BOWDataset(val_bow, idx2token), len(idx2token), id2token,
            docs_train, cv)


def get_bag_of_words(data: Sequence[Sequence[int]], min_length: int) -> sp.csr_matrix:
    """Token-id lists -> count matrix with ``min_length`` columns."""
    rows, cols = [], []
    for i, doc in enumerate(data):
        rows.extend([i] * len(doc))
        cols.extend(doc)
    m = sp.coo_matrix((np.ones(len(rows), dtype=np.float32), (rows, cols)), shape=(len(data), min_length))
    return m.tocsr()


def bert_embeddings_from_list(texts: Sequence[str], sbert_model_to_load: str, batch_size: int = 200,
                              max_seq_length: Optional[int] = None) -> np.ndarray:
    """Sentence-BERT document embeddings.  ``sentence_transformers`` is optional (not installed on the
    target boxes; the reference has the import commented out as well): pre-computed embeddings are the
    supported input."""
    try:
        from sentence_transformers import SentenceTransformer
    except Exception as ex:  # pragma: no cover - optional dependency
        raise RuntimeError("sentence_transformers is not available; pass pre-computed embeddings") from ex
    model = SentenceTransformer(sbert_model_to_load)
    if max_seq_length is not None:
        model.max_seq_length = max_seq_length
    return np.array(model.encode(list(texts), show_progress_bar=True, batch_size=batch_size))


def check_max_local_length(max_seq_length: Optional[int], texts: Sequence[str]) -> None:
    """Warns when the longest document (in whitespace tokens) exceeds what the embedding model keeps
    (utils/data_preparation.py:57-62)."""
    import warnings

    if max_seq_length is None or not len(texts):
        return
    longest = max(len(t.split()) for t in texts)
    if longest > max_seq_length:
        warnings.warn(f"the longest document in your collection has {longest} words, the model instead "
                      f"truncates to {max_seq_length} tokens.")


def bert_embeddings_from_file(text_file: str, sbert_model_to_load: str, batch_size: int = 32,
                              max_seq_length: Optional[int] = None) -> np.ndarray:
    """Sentence-BERT embeddings of a text file with one document per line (utils/data_preparation.py:25-40)."""
    with open(text_file, encoding="utf-8") as f:
        texts = f.readlines()
    check_max_local_length(max_seq_length, texts)
    return bert_embeddings_from_list(texts, sbert_model_to_load, batch_size=batch_size,
                                     max_seq_length=max_seq_length)


class TopicModelDataPreparation:
    """Builds ``CTMDataset`` objects from text (fit / transform) or from ready-made matrices (load)."""

    def __init__(self, contextualized_model: Optional[str] = None, show_warning: bool = True,
                 max_seq_length: int = 128) -> None:
        self.contextualized_model = contextualized_model
        self.vocab: List[str] = []
        self.id2token = {}
        self.vectorizer: Optional[CountVectorizer] = None
        self.label_encoder: Optional[OneHotEncoder] = None
        self.show_warning = show_warning
        self.max_seq_length = max_seq_length

    def load(self, contextualized_embeddings, bow_embeddings, id2token, labels=None) -> CTMDataset:
        return CTMDataset(X_contextual=contextualized_embeddings, X_bow=bow_embeddings, idx2token=id2token,
                          labels=labels)

    def _encode_labels(self, labels, fit: bool):
        if labels is None:
            return None
        arr = np.array([labels]).T
        if fit:
            self.label_encoder = OneHotEncoder()
            return self.label_encoder.fit_transform(arr)
        return self.label_encoder.transform(arr)

    def fit(self, text_for_contextual, text_for_bow, labels=None, custom_embeddings=None) -> CTMDataset:
        if custom_embeddings is not None and len(custom_embeddings) != len(text_for_bow):
            raise Exception("the number of custom embeddings must match the number of documents")
        if custom_embeddings is None and self.contextualized_model is None:
            raise Exception("A contextualized model or contextualized embeddings must be defined")
        self.vectorizer = CountVectorizer()
        train_bow = self.vectorizer.fit_transform(text_for_bow)
        self.vocab = self.vectorizer.get_feature_names_out()
        self.id2token = dict(enumerate(self.vocab))
        emb = custom_embeddings if custom_embeddings is not None else bert_embeddings_from_list(
            text_for_contextual, self.contextualized_model, max_seq_length=self.max_seq_length)
        return CTMDataset(X_contextual=emb, X_bow=train_bow, idx2token=self.id2token,
                          labels=self._encode_labels(labels, fit=True))

    def transform(self, text_for_contextual, text_for_bow=None, custom_embeddings=None, labels=None) -> CTMDataset:
        if text_for_bow is not None:
            test_bow = self.vectorizer.transform(text_for_bow)
        else:
            n = len(text_for_contextual) if text_for_contextual is not None else len(custom_embeddings)
            test_bow = sp.csr_matrix((n, len(self.vocab)), dtype=np.float32)  # dummy BoW (zero-shot inference)
        emb = custom_embeddings if custom_embeddings is not None else bert_embeddings_from_list(
            text_for_contextual, self.contextualized_model, max_seq_length=self.max_seq_length)
        return CTMDataset(X_contextual=emb, X_bow=test_bow, idx2token=self.id2token,
                          labels=self._encode_labels(labels, fit=False) if labels is not None else None)


def prepare_ctm_dataset(corpus: Sequence[str], unpreprocessed_corpus=None, custom_embeddings=None,
                        sbert_model_to_load: Optional[str] = None, val_size: float = 0.25,
                        max_features: Optional[int] = None):
    """Train/validation ``CTMDataset`` pair (seed 42).  Returns
    ``(train_dataset, val_dataset, input_size, id2token, qt, emb_train, emb_val, docs_train)``."""
    idx = np.arange(len(corpus))
    tr, va = train_test_split(idx, test_size=val_size, random_state=42)
    docs_train = [corpus[i] for i in tr]
    docs_val = [corpus[i] for i in va]
    raw = unpreprocessed_corpus if unpreprocessed_corpus is not None else corpus
    emb_train = emb_val = None
    if custom_embeddings is not None:
        custom_embeddings = np.asarray(custom_embeddings, dtype=np.float32)
        emb_train, emb_val = custom_embeddings[tr], custom_embeddings[va]
    qt = TopicModelDataPreparation(sbert_model_to_load)
    qt.vectorizer = CountVectorizer(input="content", lowercase=True, stop_words="english", binary=False,
                                    max_features=max_features)
    train_bow = qt.vectorizer.fit_transform(docs_train)
    val_bow = qt.vectorizer.transform(docs_val)
    qt.vocab = qt.vectorizer.get_feature_names_out()
    qt.id2token = dict(enumerate(qt.vocab))
    if emb_train is None:
        emb_train = bert_embeddings_from_list([raw[i] for i in tr], sbert_model_to_load)
        emb_val = bert_embeddings_from_list([raw[i] for i in va], sbert_model_to_load)
    train = qt.load(emb_train, train_bow, qt.id2token)
    val = qt.load(emb_val, val_bow, qt.id2token)
    return train, val, len(qt.vocab), qt.id2token, qt, emb_train, emb_val, docs_train


def prepare_hold_out_dataset(hold_out_corpus: Sequence[str], qt: TopicModelDataPreparation,
                             unpreprocessed_ho_corpus=None, embeddings_ho=None) -> CTMDataset:
    raw = unpreprocessed_ho_corpus if unpreprocessed_ho_corpus is not None else hold_out_corpus
    return qt.transform(text_for_contextual=raw, text_for_bow=hold_out_corpus, custom_embeddings=embeddings_ho)


_STOPWORDS_EN = None


def _english_stopwords():
    global _STOPWORDS_EN
    if _STOPWORDS_EN is None:
        try:
            from nltk.corpus import stopwords as nltk_sw

            _STOPWORDS_EN = set(nltk_sw.words("english"))
        except Exception:
            from sklearn.feature_extraction.text import ENGLISH_STOP_WORDS

            _STOPWORDS_EN = set(ENGLISH_STOP_WORDS)
    return _STOPWORDS_EN


class WhiteSpacePreprocessing:
    """Lower-cases, strips punctuation / digits-only tokens / stop words, keeps the ``vocabulary_size`` most
    frequent terms.  ``preprocess()`` returns ``(preprocessed_docs, unpreprocessed_docs, vocabulary)`` with
    empty documents removed from both lists."""

    def __init__(self, documents: Sequence[str], stopwords_language: str = "english", vocabulary_size: int = 2000,
                 stopwords: Optional[Sequence[str]] = None) -> None:
        self.documents = list(documents)
        self.stopwords = set(stopwords) if stopwords is not None else _english_stopwords()
        self.vocabulary_size = vocabulary_size

    def preprocess(self):
        table = str.maketrans(string.punctuation, " " * len(string.punctuation))
        docs = [d.lower().translate(table) for d in self.documents]
        docs = [" ".join(w for w in d.split() if len(w) > 0 and w not in self.stopwords and not w.isdigit())
                for d in docs]
        cv = CountVectorizer(max_features=self.vocabulary_size, token_pattern=r"\b[a-zA-Z]{2,}\b")
        cv.fit(docs)
        vocab = set(cv.get_feature_names_out())
        pre = [" ".join(w for w in d.split() if w in vocab) for d in docs]
        kept = [(p, u) for p, u in zip(pre, self.documents) if len(p) > 0]
        return [p for p, _ in kept], [u for _, u in kept], list(vocab)
