"""Evaluation path around the decoder: what `run_evaluation` (train.py:198-347) and the inference
entry (generate_caption.py:76-102) do with the hot path's outputs.

  * greedy decode: `torch.max(preds, dim=2)[1]` (train.py:272,297) on the device (`metrics.greedy_tokens`);
  * ids -> words with the reference's stopping rules (`vanilla_decode_caption` train.py:280-289,
    `bert_decode_caption` train.py:249-259; `caption_to_words` generate_caption.py:92-102);
  * BLEU-1..4 with the reference's weights (train.py:330-333).  The reference calls `nltk.translate.bleu_score.corpus_bleu`;
    nltk is a third-party dependency absent here (`requirements.txt`, unpinned), so its published algorithm is restated:
    Papineni et al. 2002 corpus BLEU = BP * exp(sum_n w_n log p_n) with clipped n-gram counts summed over the corpus,
    the closest reference length per hypothesis, and nltk's `SmoothingFunction.method0` behaviour for a zero n-gram overlap
    (p_n -> sys.float_info.min, i.e. a score of ~0) -- pinned by nltk's own documented doctest values in tests/test_evaluate.py;
  * `Evaluator`: the per-batch loop body of train.py:214-300 on top of `Decoder.forward_loss` (loss + logits in one call),
    `metrics.logit_rank` (top-1 / top-5 in one pass) and `metrics.caption_token_count`, with ONE device->host transfer per
    batch (the reference does five `.item()` syncs and two `.tolist()`).
String handling stays on the host, as upstream; everything numeric runs in the CUDA library.
"""
from __future__ import annotations

import math
import sys
from collections import Counter
from fractions import Fraction

import torch


class AverageMeter(object):
    """utils.py:4-20."""

    def __init__(self):
        self.reset()

    def reset(self):
        self.val = self.avg = self.sum = self.count = 0

    def update(self, val, n=1):
        self.val = val
        self.sum += val * n
        self.count += n
        self.avg = self.sum / self.count


# ---- ids -> words --------------------------------------------------------------------------------------------------
def vanilla_decode_caption(caption, word_dict, token_dict=None):
    """train.py:280-289: stop at <eos>, drop <start> / <pad>."""
    if token_dict is None:
        token_dict = {idx: word for word, idx in word_dict.items()}
    sentence = []
    for word_idx in caption:
        if word_idx == word_dict["<eos>"]:
            break
        if word_idx not in (word_dict["<start>"], word_dict["<pad>"]):
            sentence.append(token_dict[word_idx])
    return sentence


def bert_decode_caption(caption, tokenizer):
    """train.py:249-259: stop at [SEP], drop [CLS] / [PAD], re-join word pieces."""
    tokens = tokenizer.convert_ids_to_tokens(caption)
    sentence = []
    for token in tokens:
        if token == "[SEP]":
            break
        if token not in ("[CLS]", "[PAD]"):
            sentence.append(token)
    return tokenizer.convert_tokens_to_string(sentence).split()


def caption_to_words(sentence, tokenizer, bert=False):
    """generate_caption.py:92-102: the words of a beam-search result; the plain vocabulary keeps <eos> as last word."""
    if bert:
        return tokenizer.decode(sentence, skip_special_tokens=False).split()
    token_dict = {idx: word for word, idx in tokenizer.items()}
    words = []
    for word_idx in sentence:
        words.append(token_dict[word_idx])
        if word_idx == tokenizer["<eos>"]:
            break
    return words


# ---- BLEU (nltk.translate.bleu_score.corpus_bleu restated) -------------------------------------------------------------
def _ngrams(seq, n):
    return [tuple(seq[i:i + n]) for i in range(len(seq) - n + 1)]


def modified_precision(references, hypothesis, n):
    """Clipped n-gram precision of one hypothesis: (numerator, denominator)."""
    counts = Counter(_ngrams(hypothesis, n)) if len(hypothesis) >= n else Counter()
    max_counts = {}
    for reference in references:
        ref_counts = Counter(_ngrams(reference, n)) if len(reference) >= n else Counter()
        for ng in counts:
            max_counts[ng] = max(max_counts.get(ng, 0), ref_counts[ng])
    clipped = {ng: min(c, max_counts[ng]) for ng, c in counts.items()}
    return sum(clipped.values()), max(1, sum(counts.values()))


def closest_ref_length(references, hyp_len):
    return min((len(r) for r in references), key=lambda ref_len: (abs(ref_len - hyp_len), ref_len))


def brevity_penalty(closest_ref_len, hyp_len):
    if hyp_len > closest_ref_len:
        return 1.0
    if hyp_len == 0:
        return 0.0
    return math.exp(1 - closest_ref_len / hyp_len)


def corpus_bleu(list_of_references, hypotheses, weights=(0.25, 0.25, 0.25, 0.25)):
    """Corpus-level BLEU as nltk computes it with default arguments (no smoothing, no auto-reweigh)."""
    assert len(list_of_references) == len(hypotheses), "The number of hypotheses and their reference(s) should be the same"
    max_n = len(weights)
    num, den = Counter(), Counter()
    hyp_lengths = ref_lengths = 0
    for references, hypothesis in zip(list_of_references, hypotheses):
        for i in range(1, max_n + 1):
            a, b = modified_precision(references, hypothesis, i)
            num[i] += a
            den[i] += b
        hyp_len = len(hypothesis)
        hyp_lengths += hyp_len
        ref_lengths += closest_ref_length(references, hyp_len)
    bp = brevity_penalty(ref_lengths, hyp_lengths)
    if num[1] == 0:
        return 0
    # method0: no smoothing; an order without any match contributes log(float_info.min) (nltk warns and does the same)
    p_n = [Fraction(num[i], den[i]) if num[i] != 0 else sys.float_info.min for i in range(1, max_n + 1)]
    s = math.fsum(w * math.log(p) for w, p in zip(weights, p_n) if p > 0)
    return bp * math.exp(s)


def sentence_bleu(references, hypothesis, weights=(0.25, 0.25, 0.25, 0.25)):
    return corpus_bleu([references], [hypothesis], weights)


BLEU_WEIGHTS = {"bleu1": (1, 0, 0, 0), "bleu2": (0.5, 0.5, 0, 0), "bleu3": (0.33, 0.33, 0.33, 0),        # train.py:330-332
                "bleu4": (0.25, 0.25, 0.25, 0.25)}                                                         # train.py:333


# ---- the evaluation loop body ----------------------------------------------------------------------------------------
class Evaluator:
    """Accumulates what `run_evaluation` accumulates over the batches of a data loader (train.py:202-300)."""

    def __init__(self, decoder, word_dict=None, tokenizer=None, alpha_c=1.0):
        self.decoder, self.alpha_c = decoder, alpha_c
        self.bert = bool(decoder.use_bert)
        self.word_dict, self.tokenizer = word_dict, tokenizer if tokenizer is not None else getattr(decoder, "tokenizer", None)
        if not self.bert:
            if word_dict is None:
                raise ValueError("a plain-vocabulary decoder needs word_dict")
            self.token_dict = {idx: word for word, idx in word_dict.items()}
            self.padding_idx = word_dict["<pad>"]
            self.skip = [word_dict["<pad>"], word_dict["<start>"], word_dict["<eos>"]]                   # train.py:242
        else:
            self.padding_idx = self.tokenizer.pad_token_id
            self.skip = [self.tokenizer.pad_token_id, self.tokenizer.cls_token_id, self.tokenizer.sep_token_id]   # train.py:240
        self.losses, self.top1, self.top5 = AverageMeter(), AverageMeter(), AverageMeter()
        self.decoded_captions, self.decoded_all_captions, self.decoded_hypotheses = [], [], []

    def _decode(self, caption):
        if self.bert:
            return bert_decode_caption(caption, self.tokenizer)
        return vanilla_decode_caption(caption, self.word_dict, self.token_dict)

    @torch.no_grad()
    def update(self, img_features, captions, all_captions=None):
        """One batch: train.py:216-300.  `all_captions` (B, n_refs, Tcap) as the dataset yields them (dataset.py:45-52);
        None = the assigned caption is the only reference."""
        from . import metrics
        dec = self.decoder
        was_training = dec.training
        dec.eval()                                                                 # train.py:199-200
        try:
            captions = captions.to(img_features.device)
            loss, preds, alphas = dec.forward_loss(img_features, captions, self.alpha_c)      # train.py:217,226-232
            targets = captions[:, 1:]
            rank = metrics.logit_rank(preds, targets)
            counts = metrics.accuracy_counts(rank, targets, (1, 5), self.padding_idx)         # train.py:222-223
            ntok = metrics.caption_token_count(captions, self.skip)                           # train.py:239-242
            pred_captions = metrics.greedy_tokens(preds)                                      # train.py:272,297
            # one transfer: [loss, correct@1, correct@5, non-pad targets, caption tokens] and the greedy ids
            stats = torch.cat((loss.reshape(1).float(), counts.float(), ntok.float())).cpu().tolist()
            pred_list = pred_captions.cpu().tolist()
        finally:
            dec.train(was_training)
        n = int(stats[4])
        acc1 = stats[1] * 100.0 / stats[3] if stats[3] > 0 else 0
        acc5 = stats[2] * 100.0 / stats[3] if stats[3] > 0 else 0
        self.losses.update(stats[0], n)
        self.top1.update(acc1, n)
        self.top5.update(acc5, n)
        cap_list = captions.cpu().tolist()
        for caption in cap_list:
            self.decoded_captions.append(self._decode(caption))
        if all_captions is None:
            for caption in cap_list:
                self.decoded_all_captions.append([self._decode(caption)])
        else:
            for cap_set in (all_captions.tolist() if torch.is_tensor(all_captions) else all_captions):
                self.decoded_all_captions.append([self._decode(c) for c in cap_set])
        for pred_caption in pred_list:
            self.decoded_hypotheses.append(self._decode(pred_caption))
        return {"loss": stats[0], "top1": acc1, "top5": acc5, "tokens": n}

    def bleu(self):
        """train.py:330-333."""
        return {k: corpus_bleu(self.decoded_all_captions, self.decoded_hypotheses, weights=w) for k, w in BLEU_WEIGHTS.items()}

    def summary(self):
        out = {"loss": self.losses.avg, "top1_acc": self.top1.avg, "top5_acc": self.top5.avg}
        out.update(self.bleu())
        return out


def generate_captions(decoder, img_features, beam_size=3, tokenizer=None, bert=None):
    """Batched inference entry: generate_caption.py:88-102 for N images at once.  img_features (N, L, D), one row per image.
    Returns a list of (words, sentence ids, alpha) -- `words` is None when no tokenizer / word_dict is given."""
    bert = bool(decoder.use_bert) if bert is None else bert
    out = []
    for sentence, alpha in decoder.caption_batch(img_features, beam_size):
        words = caption_to_words(sentence, tokenizer, bert) if tokenizer is not None else None
        out.append((words, sentence, alpha))
    return out
