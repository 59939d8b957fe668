"""``SiameseGCNTNMSE``: the concrete Siamese model behind the reference's model contract
(model/Siamese/models.py:6-131 + model_mse.py:9-168), running on libsgnn.

What the reference does with 12*B placeholders and an unrolled TF graph is here:
  * every graph of the data set is packed once and kept RESIDENT in HBM (train | valid | test
    lists concatenated into one block-diagonal CSR batch);
  * a "feed dict" is just the index pairs into that batch plus the B target distances;
  * ``opt_op`` / ``loss`` / ``pred_sim_without_act()`` are tokens that ``Session.run`` maps to
    ``sgnn_train_step`` + ``sgnn_apply_adam`` / ``sgnn_eval_pairs``.
"""
import numpy as np

from .engine import SgnnEngine
from .layers_factory import create_activation, parse_layers
from .native import PackedBatch
from .similarity import create_sim_kernel


class Token(object):
    def __init__(self, name):
        self.name = name

    def __repr__(self):
        return '<{}>'.format(self.name)


class SiameseFeed(dict):
    """The packed-batch equivalent of the reference's feed dict (model_mse.py:52-94)."""

    def __init__(self, tvt, g1, g2, dists=None, norm_dists=None, pairs=None):
        super().__init__(tvt=tvt, g1=g1, g2=g2, dists=dists, norm_dists=norm_dists)
        self.tvt, self.g1, self.g2 = tvt, g1, g2
        self.dists, self.norm_dists, self.pairs = dists, norm_dists, pairs


class SiameseGCNTNMSE(object):
    def __init__(self, input_dim, **overrides):
        from .config import FLAGS, ensure_parsed, layer_specs
        ensure_parsed()
        get = lambda k: overrides[k] if k in overrides else getattr(FLAGS, k)  # noqa: E731
        assert get('loss_func') == 'mse'                                    # model_mse.py:11
        self.input_dim = input_dim
        self.batch_size = get('batch_size')
        self.dist_norm = get('dist_norm')
        self.label_mode = get('label_mode')
        self.yeta = get('yeta')
        specs = overrides.get('layer_specs') or layer_specs()
        self.layer_specs = list(specs)
        self.layers = parse_layers(specs, input_dim)                        # models.py:41-42
        assert len(self.layers) > 0
        self.sim_kernel = create_sim_kernel(get('sim_kernel'), self.yeta)   # model_mse.py:38
        self.final_act_name = get('final_act')
        self.final_act_np = create_activation(self.final_act_name, self.sim_kernel, use_tf=False)
        if self.final_act_name == 'sim_kernel' and get('sim_kernel') != 'gaussian':
            raise RuntimeError('final_act=sim_kernel on the device path needs sim_kernel=gaussian')
        self.engine = SgnnEngine(
            self.layers, input_dim, final_act=self.final_act_name, yeta=self.yeta, ntn_mode=get('ntn_mode'),
            loss_mode=get('loss_mode'), weight_decay=get('weight_decay'),
            learning_rate=get('learning_rate'), dropout=get('dropout'), precision=get('precision'),
            device=get('device'))
        self.engine.init_params(get('seed'))                                # glorot / zeros, inits.py
        self.engine.set_dropout_stream(get('seed'))                         # mask stream of FLAGS.dropout
        self.loss = Token('loss')
        self.opt_op = Token('opt_op')
        self.test_output = Token('test_output')
        self._resident = None
        print('Model built')

    # ---- reference model contract --------------------------------------------------------
    def pred_sim_without_act(self):
        return self.test_output                                             # models.py:90-91

    def apply_final_act_np(self, score):
        return self.final_act_np(score)                                     # model_mse.py:49-50

    def vars(self):
        """{'<layer>_<kind>/<var>': ndarray} in creation order (layers.py:29)."""
        flat = self.engine.get_params()
        out = {}
        for p in self.engine.param_table():
            n = int(np.prod(p['shape']))
            key = '{}_{}/{}'.format(p['layer'], self.layers[p['layer']]['kind'].lower(), p['name'])
            out[key] = flat[p['offset']:p['offset'] + n].reshape(p['shape'])
        return out

    def save(self, path):
        """Flat-blob checkpoint (the reference's tf.train.Saver path, models.py:118-131, is dead code)."""
        np.save(path, self.engine.get_params())

    def load(self, path):
        self.engine.set_params(np.load(path))

    # ---- data residency ------------------------------------------------------------------
    @staticmethod
    def _offsets(data):
        """Global graph index of each list inside the concatenated resident batch."""
        offsets, o = {}, 0
        for l in (data.train_data, data.valid_data, data.test_data):
            offsets[id(l.batch)] = o
            o += l.batch.n_graphs
        return offsets

    def _ensure_resident(self, data):
        if self._resident is not None and self._resident[0] is data:
            return self._resident
        lists = [data.train_data, data.valid_data, data.test_data]
        host = PackedBatch.concat([l.batch for l in lists])
        dev = host.to(self.engine.device)
        self._resident = (data, dev, self._offsets(data), host)
        return self._resident

    def _gid(self, offsets, mg):
        return offsets[id(mg.batch)] + mg.index

    def get_feed_dict(self, data, dist_calculator, tvt, test_id, train_id):
        offsets = self._offsets(data)
        self._data = data
        B = self.batch_size
        if tvt == 'train' or tvt == 'val':
            assert test_id is None and train_id is None
            pairs = [data.get_graph_pair(tvt) for _ in range(B)]            # model_mse.py:58-59
        else:
            assert tvt == 'test'
            pairs = [(data.test_data.get_graph(test_id), data.get_orig_train_graph(train_id))]
        g1 = np.array([self._gid(offsets, a) for a, _ in pairs], np.int32)
        g2 = np.array([self._gid(offsets, b) for _, b in pairs], np.int32)
        dists = norm_dists = None
        if tvt != 'test':
            if self.label_mode == 'reference':
                # model_mse.py:82-92 (F7): for EVERY fed pair the labels are re-drawn from B fresh
                # sampler pairs, each pass overwriting the previous one; only the last pass
                # survives, but the sampler advances B*B times.
                for _ in range(B * (B - 1)):
                    data.get_graph_pair(tvt)
                label_pairs = [data.get_graph_pair(tvt) for _ in range(B)]
            else:
                label_pairs = pairs
            dists, norm_dists = np.zeros((B, 1)), np.zeros((B, 1))
            for i, (a, b) in enumerate(label_pairs):
                dists[i], norm_dists[i] = data.get_dist(a.get_nxgraph(), b.get_nxgraph(), dist_calculator)
        return SiameseFeed(tvt, g1, g2, dists, norm_dists, pairs)

    # ---- execution -----------------------------------------------------------------------
    def targets(self, feed):
        d = feed.norm_dists if self.dist_norm else feed.dists               # model_mse.py:146
        return self.sim_kernel.dist_to_sim_np(d).astype(np.float32).ravel()

    def run(self, objs, feed):
        """Session.run for this model: returns one value per requested token."""
        import torch
        eng = self.engine
        _, dev, _, _ = self._ensure_resident(self._data)
        g1 = torch.as_tensor(feed.g1, device=eng.device)
        g2 = torch.as_tensor(feed.g2, device=eng.device)
        names = [o.name if isinstance(o, Token) else None for o in objs]
        out = [None] * len(objs)
        if 'opt_op' in names:
            t = self.targets(feed)
            td = t.astype(np.float64)
            loss = eng.train_step(dev, g1, g2, torch.as_tensor(t, device=eng.device), len(t),
                                  float(td.sum()), float((td * td).sum()))
            eng.apply_adam(1.0, loss)
            l = loss.cpu().numpy()
            total = np.float32(l[0] + l[1])                                 # mse part + weight decay part
            for i, nm in enumerate(names):
                if nm == 'loss':
                    out[i] = total
        elif 'loss' in names:
            t = self.targets(feed)
            loss, _, _ = eng.eval_pairs(dev, g1, g2, torch.as_tensor(t, device=eng.device))
            wd = 0.5 * eng.cfg.weight_decay * float(np.sum(np.square(eng.get_params().astype(np.float64))))
            out[names.index('loss')] = np.float32(float(loss[0]) + wd)
        if 'test_output' in names:
            _, raw, _ = eng.eval_pairs(dev, g1, g2)
            out[names.index('test_output')] = [float(v) for v in raw.cpu().numpy()]  # list of B scalars
        return out
