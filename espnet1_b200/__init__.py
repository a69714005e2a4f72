"""espnet1_b200 - B200-native (sm_100a) implementation of ESPnet's CTC hot path.

Drop-in classes (same names / signatures as the reference):
  espnet1_b200.espnet2.asr.ctc.CTC
  espnet1_b200.espnet.nets.pytorch_backend.ctc.CTC, ctc_for
  espnet1_b200.espnet.nets.ctc_prefix_score.CTCPrefixScoreTH
  espnet1_b200.espnet.nets.scorers.ctc.CTCPrefixScorer
All arithmetic runs in libb200ctc.so (C ABI: include/b200ctc.h); there is no
CPU or PyTorch fallback.  See DESIGN.md and INTEGRATION.md.
"""
__version__ = "0.1.0"
