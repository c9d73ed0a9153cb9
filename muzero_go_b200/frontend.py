"""Checkpoint compatibility and the `play.py` front end -- SURVEY §8(f) rank 4.

  print_board / action_to_coord / coord_to_action     /root/reference/play.py:248-278 (same text, same quirks:
                                                       'pass' is matched before stripping, a malformed row raises
                                                       ValueError("Invalid coordinate format"))
  checkpoints                                          main.py:750-762 (torch.save(state_dict)), play.py:288-295,
                                                       self_play.py:404-412 (torch.load + load_state_dict): the
                                                       state_dict key names and shapes are the reference's
  play()                                               play.py:281-348, the interactive loop, with the terminal
                                                       I/O injectable
"""
import io

import torch


def format_board(obs, board_size):
    """The text play.py:248-259 prints: 'B' / 'W' / '.', column letters on top, 1-based row numbers."""
    out = io.StringIO()
    board = [["." for _ in range(board_size)] for _ in range(board_size)]
    for i in range(board_size):
        for j in range(board_size):
            if obs[0, i, j] == 1:
                board[i][j] = "B"
            elif obs[1, i, j] == 1:
                board[i][j] = "W"
    out.write("   " + " ".join(chr(ord("A") + j) for j in range(board_size)) + "\n")
    for i, row in enumerate(board):
        out.write(f"{i + 1:2d} " + " ".join(row) + "\n")
    return out.getvalue()


def print_board(obs, board_size, print_fn=print):
    print_fn(format_board(obs, board_size), end="")


def action_to_coord(action, board_size):
    """play.py:261-266: a = row * N + col -> 'A1'-style (column letter, 1-based row); N*N -> 'pass'."""
    if action == board_size * board_size:
        return "pass"
    return f"{chr(ord('A') + action % board_size)}{action // board_size + 1}"


def coord_to_action(coord, board_size):
    """play.py:268-278."""
    if coord.lower() == "pass":
        return board_size * board_size
    coord = coord.strip().upper()
    col_letter = coord[0]
    try:
        row = int(coord[1:]) - 1
    except Exception:
        raise ValueError("Invalid coordinate format")
    return row * board_size + (ord(col_letter) - ord("A"))


def save_checkpoint(net, path):
    """main.py:752-753, :760-761: the bare state_dict."""
    torch.save(net.state_dict(), path)


def load_checkpoint(net, path, map_location=None):
    """play.py:290-291 / self_play.py:406-408: works for files the reference wrote and for `save_checkpoint`'s."""
    state = torch.load(path, map_location=map_location if map_location is not None else net.prediction.pass_logit.device)
    net.load_state_dict(state)
    return net


def play(agent, env, board_size, human_is_black, input_fn=input, print_fn=print):
    """play.py:305-346 with the agent, environment and terminal passed in.  Returns (outcome, reward)."""
    observation = env.reset()
    if isinstance(observation, tuple):
        observation = observation[0]
    done, move_count, reward = False, 0, 0
    max_moves = board_size * board_size * 2
    human_turn = human_is_black
    while not done and move_count < max_moves:
        print_board(observation, board_size, print_fn)
        if human_turn:
            move = input_fn("Your move (e.g., A1 or 'pass'): ")
            try:
                action = coord_to_action(move, board_size)
            except Exception:
                print_fn("Invalid move format. Try again.")
                continue
        else:
            action = agent.select_action(observation)
            print_fn(f"Agent move: {action_to_coord(action, board_size)}")
        result = env.step(action)
        if len(result) == 5:
            observation, reward, done, truncated, _ = result
            done = done or truncated
        else:
            observation, reward, done, _ = result
        if isinstance(observation, tuple):
            observation = observation[0]
        human_turn = not human_turn
        move_count += 1
    if not done:
        print_fn(f"Reached max moves {move_count}, forcing game end.")
    print_board(observation, board_size, print_fn)
    print_fn("Game over!")
    if reward > 0:
        outcome = "win" if human_is_black else "loss"
    elif reward < 0:
        outcome = "loss" if human_is_black else "win"
    else:
        outcome = "draw"
    print_fn(f"Final outcome: {outcome}. Final reward: {reward}")
    return outcome, reward
