"""MuZeroAgent with the reference constructors' signature
(/root/reference/main.py:371-376, play.py:219-245, self_play.py:347-412): owns a MuZeroNet and picks moves
from the search's visit counts.  Search and network run on the GPU through libmzgo.so.

select_action(obs)               -> int                                  (play.py:226-245, greedy)
select_action(obs, temperature)  -> (action, policy_target, root_value)  (self_play.py:357-402; the visit
    counts are read from the child nodes -- the reference's returned vector is always zero, a defect
    documented in SURVEY.md section 0, which would make it play uniformly at random)
"""
import random
import sys

import numpy as np
import torch

from .mcts import MCTS, valid_mask_from_observation
from .networks import MuZeroNet
from .tree import SearchConfig


class MuZeroAgent:
    def __init__(self, board_size, latent_dim, env_action_size, num_simulations, max_action_size=None,
                 precision="fp32", config=None, c_puct=None, device="cuda", variant="main"):
        self.board_size = board_size
        self.action_size = env_action_size
        if max_action_size is None:
            max_action_size = board_size * board_size + 1          # main.py:30 / self_play.py:353
        self.net = MuZeroNet(latent_dim, max_action_size, board_size=board_size, precision=precision).to(device)
        self.net.eval()
        self.mcts_simulations = num_simulations
        self.variant = variant
        self.config = config if config is not None else SearchConfig.for_variant(variant)
        self.c_puct = c_puct
        self._mcts = None

    def _search(self, observation):
        if self._mcts is None:
            self._mcts = MCTS(self.net, self.action_size, self.mcts_simulations, c_puct=self.c_puct,
                              config=self.config, variant=self.variant)
        out = self._mcts.run(observation)
        root = out[0] if isinstance(out, tuple) else out
        valid_mask = valid_mask_from_observation(observation, self.config.pass_epsilon)
        visit_counts = np.array([child["node"].visit_count if (child["node"] is not None and valid_mask[a] > 0) else 0
                                 for a, child in root.children.items()])
        return root, visit_counts, valid_mask

    def select_action(self, observation, temperature=None):
        root, visit_counts, valid_mask = self._search(observation)
        if temperature is None:                                    # play.py / main.py: greedy, returns the action
            if visit_counts.sum() > 0:
                return int(np.argmax(visit_counts))
            return int(random.choice(np.nonzero(valid_mask)[0]))
        visit_counts = visit_counts * valid_mask                   # self_play.py:376
        total = visit_counts.sum()
        policy_target = visit_counts / total if total > 0 else valid_mask / valid_mask.sum()
        if temperature == 0:
            if total > 0:
                action = int(np.argmax(visit_counts))
            else:
                action = int(random.choice(np.where(valid_mask > 0)[0]))
        else:
            vt = visit_counts ** (1.0 / temperature) * valid_mask
            probs = vt / vt.sum() if vt.sum() > 0 else valid_mask / valid_mask.sum()
            action = int(np.random.choice(self.action_size, p=probs))
        return action, policy_target, root.value()

    def load_weights(self, weight_file):
        try:
            state_dict = torch.load(weight_file, map_location=self.net.prediction.pass_logit.device)
            self.net.load_state_dict(state_dict)
            self.net.eval()
            print(f"Loaded weights from {weight_file}")
        except Exception as e:
            print(f"Error loading weight file '{weight_file}': {e}")
            sys.exit(1)
