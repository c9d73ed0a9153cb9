"""Oracle: restatement of the reference MCTS (TEST INFRASTRUCTURE, see oracle/__init__.py).

Restates /root/reference/main.py:247-368 (MCTS.run / select_leaf / backpropagate,
the variant whose visit counts are consumed at main.py:666-672) and, through
`c_puct`, play.py:104-216 (same algorithm, c_puct=10.0, play.py:116).  The
arithmetic keeps the reference's dtypes on purpose (numpy >= 2 / NEP 50):

  * root priors  float64  (float32 softmax * float64 mask -> float32, pairwise
    float32 sum, float32 divide, then mixed with float64 Dirichlet noise;
    main.py:259-282, :62-64);
  * child priors float32  (main.py:298-304), zeroed where the ROOT mask is 0;
  * PUCT score   float64: q_norm + (c*P)*sqrt(N_parent+1)/(1+n) with c*P
    evaluated in P's dtype (main.py:349-355);
  * backup       float(reward) + discount*float(value), added unchanged to
    every node on the path (main.py:311, :365-368).

Tree storage is structure-of-arrays (not the reference's dict-of-dicts) so this
file is an independent restatement; it is pinned to the real reference by
tests/golden/mcts_*.npz (oracle/make_golden.py).

Randomness is injected: `noise` replaces np.random.dirichlet (main.py:63) and
`choice_stream[i]` replaces the i-th random.choice(unexpanded) call
(main.py:324) as unexpanded[stream[i] % len(unexpanded)].
"""
import numpy as np

INVD_CHNL = 3


def valid_mask_from_obs(obs, pass_epsilon):
    """main.py:261-267 -> float64 [A]; board entries 0/1, pass entry pass_epsilon or 1."""
    valid_board = (np.asarray(obs)[INVD_CHNL].flatten() == 0).astype(np.float32)
    pass_prior = pass_epsilon if valid_board.sum() > 0 else 1.0
    return np.concatenate([valid_board, np.array([pass_prior])])


def root_priors(softmax_f32, valid_mask, noise, epsilon):
    """main.py:269-282 -> float64 [A] root edge priors."""
    policy = np.array(softmax_f32, dtype=np.float32, copy=True)
    policy *= valid_mask
    total = policy.sum()
    if total > 0:
        policy /= total
    mixed = (1 - epsilon) * policy + epsilon * np.asarray(noise, dtype=np.float64)
    return np.where(valid_mask > 0, mixed, 0.0)


class TreeRef:
    """SoA search tree with the reference's select / expand / backup arithmetic."""

    def __init__(self, action_size, c_puct=2, discount=0.99):
        self.A = action_size
        self.c_puct = c_puct
        self.discount = discount
        self.prior = []       # per node: [A] array (float64 for the root, float32 below)
        self.child = []       # per node: [A] int, -1 = edge not expanded
        self.visit = []       # per node: int
        self.value_sum = []   # per node: float
        self.latent = []      # per node: opaque handle owned by the network callbacks
        self.dropped = 0      # simulations that found no edge with prior > 0 (main.py:359-363)
        self.depths = []      # leaf depth of every completed simulation

    def add_node(self, prior, latent):
        self.prior.append(prior)
        self.child.append(np.full(self.A, -1, dtype=np.int64))
        self.visit.append(0)
        self.value_sum.append(0)
        self.latent.append(latent)
        return len(self.visit) - 1

    def q(self, node):
        return self.value_sum[node] / self.visit[node] if self.visit[node] > 0 else 0.0

    def select(self, choose):
        """main.py:318-363.  Returns (path, action) ; action None = nothing selectable."""
        node = 0
        path = [0]
        while True:
            pri, ch = self.prior[node], self.child[node]
            open_edges = [a for a in range(self.A) if pri[a] > 0 and ch[a] < 0]
            if open_edges:
                return path, choose(open_edges)
            live = [a for a in range(self.A) if pri[a] > 0]
            if not live:
                return path, None
            qs = [self.q(int(ch[a])) for a in live]
            lo, hi = min(qs), max(qs)
            best, best_a = -float('inf'), None
            for a, qa in zip(live, qs):
                qn = (qa - lo) / (hi - lo) if hi > lo else qa
                n = self.visit[int(ch[a])]
                u = self.c_puct * pri[a] * np.sqrt(self.visit[node] + 1) / (1 + n)
                score = qn + u
                if score > best:
                    best, best_a = score, a
            if best_a is None:      # only reachable with NaN scores
                return path, None
            node = int(ch[best_a])
            path.append(node)

    def expand_and_backup(self, path, action, child_prior_f32, root_mask, reward, value, latent):
        """main.py:300-313."""
        prior = np.where(root_mask > 0, child_prior_f32, np.float32(0)).astype(np.float32)
        new = self.add_node(prior, latent)
        self.child[path[-1]][action] = new
        backup = float(reward) + self.discount * float(value)
        for node in path + [new]:
            self.visit[node] += 1
            self.value_sum[node] += backup
        self.depths.append(len(path))

    def root_visit_counts(self, root_mask):
        """main.py:666-669."""
        ch = self.child[0]
        return np.array([self.visit[int(ch[a])] if (ch[a] >= 0 and root_mask[a] > 0) else 0
                         for a in range(self.A)], dtype=np.int64)

    def root_child_value_sums(self):
        ch = self.child[0]
        return np.array([float(self.value_sum[int(ch[a])]) if ch[a] >= 0 else 0.0
                         for a in range(self.A)], dtype=np.float64)


def run_mcts(obs, initial_fn, recurrent_fn, num_simulations, noise, choice_stream,
             c_puct=2, discount=0.99, dirichlet_epsilon=0.25, pass_epsilon=0.05):
    """main.py:255-316 with injected randomness and network callbacks.

    initial_fn(obs)               -> (latent, value_f32, softmax_f32[A])
    recurrent_fn(latent, action)  -> (latent', reward_f32, value_f32, softmax_f32[A])
    Returns (tree, root_mask).
    """
    latent, _value, soft = initial_fn(obs)
    A = len(soft)
    mask = valid_mask_from_obs(obs, pass_epsilon)
    tree = TreeRef(A, c_puct=c_puct, discount=discount)
    tree.add_node(root_priors(soft, mask, noise, dirichlet_epsilon), latent)
    calls = [0]

    def choose(seq):
        k = int(choice_stream[calls[0]]) % len(seq)
        calls[0] += 1
        return seq[k]

    for _ in range(num_simulations):
        path, action = tree.select(choose)
        if action is None:
            tree.dropped += 1
            continue
        nxt, reward, value, soft_c = recurrent_fn(tree.latent[path[-1]], action)
        tree.expand_and_backup(path, action, np.asarray(soft_c, dtype=np.float32), mask,
                               reward, value, nxt)
    return tree, mask


# ---------------------------------------------------------------------------------------------
# self_play.py variant (/root/reference/self_play.py:142-343).  Differences from main.py, all on the hot path:
#   * root priors re-masked and renormalised in float64 after the noise mix (:166-178);
#   * child priors = float32 softmax * float64 mask -> float32, float32 pairwise sum, float32 divide (:210-216);
#   * exploration term uses sqrt(max(1, N_parent)) (:297-300);
#   * a node without any prior > 0 is terminal: the simulation backs 0 up along the path (:188-190, :245-247, :268-270);
#   * backup alternates sign per level starting with +value at the new leaf (:340-343).
# At PUCT time every live child is expanded (the unexpanded branch is taken first, :283-287), so the
# "expanded children only" min/max set (:262-277) equals the set of all edges with prior > 0.
def _uniform_over(mask, like):
    """self_play.py:163, :178, :216: the rare fallback when a masked policy sums to zero."""
    ms = mask.sum()
    return np.ones_like(like) * mask / ms if ms > 0 else np.ones_like(like) / len(like)


def root_priors_selfplay(softmax_f32, valid_mask, noise, epsilon):
    """self_play.py:158-178 -> float64 [A]."""
    policy = np.array(softmax_f32, dtype=np.float32, copy=True)
    policy *= valid_mask
    total = policy.sum()
    if total > 0:
        policy /= total
    else:
        policy = _uniform_over(valid_mask, policy)
    policy = (1 - epsilon) * policy + epsilon * np.asarray(noise, dtype=np.float64)
    policy *= valid_mask
    total = policy.sum()
    if total > 0:
        policy /= total
    else:
        policy = _uniform_over(valid_mask, policy)
    return np.where(valid_mask > 0, policy, 0.0)


def child_priors_selfplay(softmax_f32, valid_mask):
    """self_play.py:210-222 -> [A] (float32 unless the zero-sum fallback promoted it to float64)."""
    policy = np.array(softmax_f32, dtype=np.float32, copy=True)
    policy *= valid_mask
    total = policy.sum()
    if total > 0:
        policy /= total
    else:
        policy = _uniform_over(valid_mask, policy)
    return np.where(valid_mask > 0, policy, policy.dtype.type(0))


class TreeRefSelfPlay(TreeRef):
    def select(self, choose):
        """self_play.py:239-331.  Returns (path, action); action None = terminal leaf."""
        node = 0
        path = [0]
        while True:
            pri, ch = self.prior[node], self.child[node]
            live = [a for a in range(self.A) if pri[a] > 0]
            if not live:                        # :245-247 / :268-270
                return path, None
            open_edges = [a for a in live if ch[a] < 0]
            if open_edges:                      # :283-287
                return path, choose(open_edges)
            qs = [self.q(int(ch[a])) for a in live]
            lo, hi = min(qs), max(qs)
            best, best_a = -float('inf'), None
            for a, qa in zip(live, qs):
                qn = (qa - lo) / (hi - lo) if hi > lo else qa
                n = self.visit[int(ch[a])]
                u = self.c_puct * pri[a] * np.sqrt(max(1, self.visit[node])) / (1 + n)
                score = qn + u
                if score > best:
                    best, best_a = score, a
            if best_a is None:                  # :315-320
                return path, None
            node = int(ch[best_a])
            path.append(node)

    def backup(self, path, value):
        """self_play.py:337-343."""
        for i, node in enumerate(reversed(path)):
            self.visit[node] += 1
            self.value_sum[node] += value * ((-1) ** i)

    def expand_and_backup(self, path, action, child_prior, root_mask, reward, value, latent):
        """self_play.py:218-230."""
        new = self.add_node(child_prior, latent)
        self.child[path[-1]][action] = new
        self.backup(path + [new], float(reward) + self.discount * float(value))
        self.depths.append(len(path))


def run_mcts_selfplay(obs, initial_fn, recurrent_fn, num_simulations, noise, choice_stream,
                      c_puct=2.5, discount=0.99, dirichlet_epsilon=0.02, pass_epsilon=0.01):
    """self_play.py:148-237 with injected randomness and network callbacks (same callbacks as run_mcts)."""
    latent, _value, soft = initial_fn(obs)
    A = len(soft)
    mask = valid_mask_from_obs(obs, pass_epsilon)
    tree = TreeRefSelfPlay(A, c_puct=c_puct, discount=discount)
    tree.add_node(root_priors_selfplay(soft, mask, noise, dirichlet_epsilon), latent)
    calls = [0]

    def choose(seq):
        k = int(choice_stream[calls[0]]) % len(seq)
        calls[0] += 1
        return seq[k]

    for _ in range(num_simulations):
        path, action = tree.select(choose)
        if action is None:
            tree.dropped += 1
            tree.backup(path, 0)                # :188-195
            continue
        nxt, reward, value, soft_c = recurrent_fn(tree.latent[path[-1]], action)
        tree.expand_and_backup(path, action, child_priors_selfplay(soft_c, mask), mask, reward, value, nxt)
    return tree, mask


def choose_action(visit_counts, root_mask, fallback_u32=0):
    """main.py:670-677 -> (action, policy_target float64[A])."""
    vc = np.asarray(visit_counts)
    if vc.sum() > 0:
        return int(np.argmax(vc)), vc / vc.sum()
    valid = np.nonzero(root_mask)[0]
    action = int(valid[int(fallback_u32) % len(valid)])
    policy = np.zeros(len(vc))
    policy[action] = 1.0
    return action, policy


def torch_net_callbacks(net):
    """Wrap a torch MuZeroNet-like module (reference or oracle) as run_mcts callbacks,
    applying torch.softmax exactly where the reference does (main.py:259, :298)."""
    import torch

    @torch.no_grad()
    def initial(obs):
        x = torch.FloatTensor(np.asarray(obs)).unsqueeze(0)
        latent, value, logits = net.initial_inference(x)
        return latent[0], value.numpy().reshape(-1)[0], torch.softmax(logits, dim=1).numpy()[0]

    @torch.no_grad()
    def recurrent(latent, action):
        nxt, reward, value, logits = net.recurrent_inference(latent.unsqueeze(0), torch.LongTensor([action]))
        return (nxt[0], reward.numpy().reshape(-1)[0], value.numpy().reshape(-1)[0],
                torch.softmax(logits, dim=1).numpy()[0])

    return initial, recurrent
