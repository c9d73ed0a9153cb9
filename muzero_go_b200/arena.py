"""Arena evaluator on the batched engine -- SURVEY §8(f) rank 3.

The reference (`SelfPlayEvaluator`, /root/reference/main.py:526-611, gating :737-748) plays `num_games` games one
after the other between the current and the best network, alternating who moves first, every move an MCTS of the
side to move (`main.py` variant: root noise included) followed by the argmax of the visit counts.  Here all games
run at once in lock step: at every ply BOTH networks search ALL games on the same positions (two
`SelfPlayEngine`s, one move each) and every game keeps the move of the network whose turn it is.  That doubles
the search work and needs nothing new on the device; games finish by GymGo's end, two passes in a row or the
move cap, exactly like the engine's self-play.

Scoring, Elo and promotion follow the reference line by line -- including its treatment of the games the best
network opens: `play_game(starting_player=1)` already returns 1 when the current network (white) wins, and
`evaluate` then inverts that result again (main.py:598-600), so in those games a LOSS or a draw of the current
network is counted as its win.  `win_rate` reproduces that (it is the number the promotion decision uses);
`true_win_rate` is the fraction of games the current network actually won.
"""
import dataclasses

import torch

from .rules import GoBoards
from .selfplay import SelfPlayEngine
from .tree import SearchConfig


@dataclasses.dataclass
class ArenaConfig:
    """main.py:43-48."""
    initial_elo: float = 1000
    elo_k: float = 32
    evaluation_num_games: int = 50
    evaluation_win_threshold: float = 0.55


def game_result(starting_player, winner):
    """What `play_game` returns (main.py:589-593): starting_player 0 = the current network plays black."""
    if starting_player == 0:
        return 1 if winner == 1 else 0
    return 1 if winner == -1 else 0


def score_games(results):
    """`evaluate`'s tally (main.py:595-602) from the per-game `play_game` results: game i has starting player
    i % 2, and the odd games' results are inverted before they are added."""
    wins = 0
    for i, r in enumerate(results):
        wins += (1 - r) if i % 2 == 1 else r
    return wins / len(results)


def elo_update(win_rate, current_elo, best_elo, cfg):
    """main.py:603-606: both ratings move only when the win rate clears the promotion threshold."""
    expected = 1 / (1 + 10 ** ((best_elo - current_elo) / 400))
    if win_rate > cfg.evaluation_win_threshold:
        current_elo += cfg.elo_k * (win_rate - expected)
        best_elo += cfg.elo_k * ((1 - win_rate) - (1 - expected))
    return current_elo, best_elo


class ArenaEvaluator:
    def __init__(self, current_net, best_net, board_size, num_simulations, num_games=None, cfg=None, search=None,
                 max_game_moves=None, komi=0.0):
        self.cfg = cfg or ArenaConfig()
        self.n, self.S = int(board_size), int(num_simulations)
        self.G = int(num_games if num_games is not None else self.cfg.evaluation_num_games)
        self.komi = float(komi)
        base = search or SearchConfig()
        # two independent RNG key spaces, as two MCTS objects drawing from one global RNG would be
        cfgs = [dataclasses.replace(base, game_base=base.game_base), dataclasses.replace(base, game_base=base.game_base + self.G)]
        self.engines = [SelfPlayEngine(net, self.n, self.G, self.S, search=c, max_game_moves=max_game_moves,
                                       restart_finished=False, traj_capacity=1, komi=komi)
                        for net, c in zip((current_net, best_net), cfgs)]
        self.max_game_moves = self.engines[0].max_game_moves
        self.device = self.engines[0].device
        self.current_elo = self.cfg.initial_elo          # main.py:532-533
        self.best_elo = self.cfg.initial_elo
        self.last = None

    @torch.no_grad()
    def play(self):
        """All games, lock step.  -> dict(winner int32 [G] (+1 black, -1 white, 0 draw or not ended: env.winner()),
        starting_player [G], actions int32 [plies, G] (-1 once a game is over), plies)."""
        G, dev = self.G, self.device
        e0, e1 = self.engines
        boards = GoBoards(G, self.n, dev).reset()
        move_count = torch.zeros(G, dtype=torch.int32, device=dev)
        pass_streak = torch.zeros(G, dtype=torch.int32, device=dev)
        starting = torch.arange(G, device=dev, dtype=torch.int32) % 2          # main.py:597
        actions = []
        for ply in range(self.max_game_moves):
            for e in (e0, e1):
                e.boards.records.copy_(boards.records)
                e.move_count.copy_(move_count)
                e.pass_streak.copy_(pass_streak)
                e.run(1)
            mine = ((starting + ply) % 2) == 0                                  # turn == 0: the current network moves
            boards.records.copy_(torch.where(mine[:, None], e0.boards.records, e1.boards.records))
            move_count = torch.where(mine, e0.move_count, e1.move_count)
            pass_streak = torch.where(mine, e0.pass_streak, e1.pass_streak)
            actions.append(torch.where(mine, e0.traj_action[0], e1.traj_action[0]))
            over = boards.done().bool() | (pass_streak >= 2) | (move_count >= self.max_game_moves)
            if bool(over.all()):
                break
        _, score_winner = boards.score(self.komi)
        winner = torch.where(boards.done().bool(), score_winner, torch.zeros_like(score_winner))
        return {"winner": winner.cpu(), "starting_player": starting.cpu(), "actions": torch.stack(actions).cpu(),
                "plies": len(actions)}

    def evaluate(self):
        """main.py:595-608 -> (win_rate, current_elo); details in `self.last`."""
        games = self.play()
        results = [game_result(int(s), int(w)) for s, w in zip(games["starting_player"], games["winner"])]
        win_rate = score_games(results)
        self.current_elo, self.best_elo = elo_update(win_rate, self.current_elo, self.best_elo, self.cfg)
        self.last = dict(games, results=results, win_rate=win_rate, true_win_rate=sum(results) / len(results),
                         promote=win_rate > self.cfg.evaluation_win_threshold)
        return win_rate, self.current_elo

    def gate(self, current_net, best_net, replay_buffer=None):
        """main.py:741-748 after an `evaluate`: promote (best <- current, new replay version) or revert."""
        if self.last["promote"]:
            best_net.load_state_dict(current_net.state_dict())
            if replay_buffer is not None:
                replay_buffer.add_version()
        else:
            current_net.load_state_dict(best_net.state_dict())
        return self.last["promote"]
