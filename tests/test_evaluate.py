"""CPU: host-side evaluation helpers against known answers.

BLEU: nltk is absent in this image, so `evaluate.corpus_bleu` (a restatement of nltk.translate.bleu_score.corpus_bleu, the
function train.py:330-333 calls) is pinned by the values nltk's own documentation publishes for its doctest corpus
(corpus_bleu 0.5920..., sentence_bleu 0.5045... / 0.7400..., their mean 0.6223...; modified precision 2/7 of Papineni et
al. 2002) and by hand-computed cases.  Decoding rules: train.py:249-259, 280-289; generate_caption.py:92-102."""
import math

import pytest

from sat_b200 import evaluate as E

hyp1 = ['It', 'is', 'a', 'guide', 'to', 'action', 'which', 'ensures', 'that', 'the', 'military', 'always', 'obeys', 'the',
        'commands', 'of', 'the', 'party']
ref1a = ['It', 'is', 'a', 'guide', 'to', 'action', 'that', 'ensures', 'that', 'the', 'military', 'will', 'forever', 'heed',
         'Party', 'commands']
ref1b = ['It', 'is', 'the', 'guiding', 'principle', 'which', 'guarantees', 'the', 'military', 'forces', 'always', 'being',
         'under', 'the', 'command', 'of', 'the', 'Party']
ref1c = ['It', 'is', 'the', 'practical', 'guide', 'for', 'the', 'army', 'always', 'to', 'heed', 'the', 'directions', 'of',
         'the', 'party']
hyp2 = ['he', 'read', 'the', 'book', 'because', 'he', 'was', 'interested', 'in', 'world', 'history']
ref2a = ['he', 'was', 'interested', 'in', 'world', 'history', 'because', 'he', 'read', 'the', 'book']


def test_bleu_matches_nltk_documented_values():
    assert E.corpus_bleu([[ref1a, ref1b, ref1c], [ref2a]], [hyp1, hyp2]) == pytest.approx(0.5920778868801042, abs=1e-12)
    s1 = E.sentence_bleu([ref1a, ref1b, ref1c], hyp1)
    s2 = E.sentence_bleu([ref2a], hyp2)
    assert s1 == pytest.approx(0.5045666840058485, abs=1e-12)
    assert (s1 + s2) / 2 == pytest.approx(0.6223247442490669, abs=1e-9)
    # Papineni's clipping example: 'the' x7 against two references -> 2/7
    assert E.modified_precision([['the', 'cat', 'is', 'on', 'the', 'mat'], ['there', 'is', 'a', 'cat', 'on', 'the', 'mat']],
                                ['the'] * 7, 1) == (2, 7)


def test_bleu_hand_computed_and_edge_cases():
    ref, hyp = ['a', 'b', 'c', 'd', 'e'], ['a', 'b', 'c', 'x', 'e']
    # unigrams 4/5, bigrams 2/4, trigrams 1/3, 4-grams 0/2 -> method0: float min for the empty order
    w = E.BLEU_WEIGHTS
    assert E.corpus_bleu([[ref]], [hyp], w["bleu1"]) == pytest.approx(0.8)
    assert E.corpus_bleu([[ref]], [hyp], w["bleu2"]) == pytest.approx(math.sqrt(0.8 * 0.5))
    assert E.corpus_bleu([[ref]], [hyp], w["bleu3"]) == pytest.approx(math.exp(0.33 * (math.log(0.8) + math.log(0.5) + math.log(1 / 3))))
    assert 0 < E.corpus_bleu([[ref]], [hyp], w["bleu4"]) < 1e-70
    assert E.corpus_bleu([[ref]], [ref]) == pytest.approx(1.0)
    assert E.corpus_bleu([[ref]], [['q', 'r']]) == 0                        # no unigram overlap
    # brevity penalty: hypothesis shorter than the closest reference
    assert E.corpus_bleu([[ref]], [['a', 'b', 'c']], (1, 0, 0, 0)) == pytest.approx(math.exp(1 - 5 / 3))
    assert E.closest_ref_length([['x'] * 4, ['x'] * 8], 6) == 4                 # tie in distance -> the shorter reference
    assert E.brevity_penalty(5, 0) == 0.0


def test_decode_rules():
    wd = {"<start>": 0, "<eos>": 1, "<unk>": 2, "<pad>": 3, "a": 4, "dog": 5, "runs": 6}
    assert E.vanilla_decode_caption([0, 4, 5, 3, 6, 1, 5, 3], wd) == ["a", "dog", "runs"]          # stops at <eos>, drops <start>/<pad>
    assert E.vanilla_decode_caption([4, 4], wd) == ["a", "a"]
    assert E.caption_to_words([0, 4, 5, 1, 6], wd) == ["<start>", "a", "dog", "<eos>"]              # generate_caption.py:97-102

    class Tok:
        vocab = {0: "[PAD]", 101: "[CLS]", 102: "[SEP]", 5: "play", 6: "##ing", 7: "ball"}

        def convert_ids_to_tokens(self, ids):
            return [self.vocab[i] for i in ids]

        def convert_tokens_to_string(self, toks):
            return " ".join(toks).replace(" ##", "")

    assert E.bert_decode_caption([101, 5, 6, 0, 7, 102, 7], Tok()) == ["playing", "ball"]


def test_average_meter():
    m = E.AverageMeter()
    m.update(2.0, 10)
    m.update(4.0, 30)
    assert m.val == 4.0 and m.count == 40 and m.avg == pytest.approx(3.5)
