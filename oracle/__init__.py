"""oracle/ — CPU restatement of the reference's attack hot path.  TEST INFRASTRUCTURE ONLY.

Nothing under image-recognition-adversarial-example-attack_b200/ imports this package; it is used
by tests/, by __graft_entry__.smoke() as the checker, and by bench.py's cpu_baseline / --impl
reference legs.  The product path has no CPU fallback.

Where the algorithm lives: the reference (picklerick422/Image-recognition-adversarial-example-attack)
is 140 lines of Python (attack.py) plus the defenses in defense_experiments.py; the arithmetic is
delegated to third-party PyTorch / torchvision (un-pinned in the reference, README.md:14-21; the
effective pin is what this image ships: torch 2.11.0, torchvision 0.26.0).  The restatement
therefore re-expresses the reference's control flow with explicit torch/numpy arithmetic
(explicit Adam, explicit 3x3 mean / quantiser, functional ResNet-50 from a state_dict) and is
pinned against outputs of the reference itself, imported from /root/reference by
oracle/gen_golden.py and committed under tests/golden/ (the reference has no tests or golden
vectors of its own).
"""
