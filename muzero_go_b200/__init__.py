"""B200-native MuZero-Go self-play hot path (drop-in for Sir-Teo/MuZero-Go's MCTS / networks / GymGo surface).

Everything that computes goes through libmzgo.so (hand-written sm_100a CUDA behind the C ABI of
include/mzgo.h).  There is no CPU fallback.
"""
from . import _lib
from ._lib import MzgoError
from .rules import GoBoards, GoEnv, govars, make
from .tree import SearchConfig, Tree
from .networks import MuZeroNet
from .mcts import MCTS, MCTSNode
from .agent import MuZeroAgent
from .selfplay import SelfPlayEngine, WaveGroup
from .replay import MultiVersionReplayBuffer, PrioritizedReplayBuffer, episodes_from_rows, load_episodes, save_episodes
from .trainer import BatchedTrainer, TrainConfig
from .arena import ArenaConfig, ArenaEvaluator
from .device_replay import DeviceReplay, train_on_device

__all__ = ["MzgoError", "GoBoards", "GoEnv", "govars", "make", "SearchConfig", "Tree", "MuZeroNet", "MCTS", "WaveGroup",
           "MCTSNode", "MuZeroAgent", "SelfPlayEngine", "MultiVersionReplayBuffer", "PrioritizedReplayBuffer",
           "episodes_from_rows", "load_episodes", "save_episodes", "BatchedTrainer", "TrainConfig", "ArenaConfig", "ArenaEvaluator", "DeviceReplay", "train_on_device"]
