"""``SiameseModelData`` (reference model/Siamese/data_siamese.py:9-113): loads the train and
test sets, splits train into train/validation by POSITION (first 1-valid_percentage, :80-90),
fits one label encoder over train+test (:31-32) and wraps the three lists in
``ModelGraphList``s.  Flags can be overridden by keyword for use without absl."""
from .graphs import ModelGraphList, NodeFeatureOneHotEncoder
from .utils import load_data


class SiameseModelData(object):
    def __init__(self, **overrides):
        from .config import FLAGS, ensure_parsed
        ensure_parsed()
        get = lambda k: overrides[k] if k in overrides else getattr(FLAGS, k)  # noqa: E731
        self.dataset = get('dataset')
        self.valid_percentage = get('valid_percentage')
        self.node_feat_name = get('node_feat_name')
        self.node_feat_encoder = get('node_feat_encoder')
        self.edge_feat_name = get('edge_feat_name')
        self.edge_feat_processor = get('edge_feat_processor')
        self.sampler = get('sampler')
        self.sample_num = get('sample_num')
        self.sampler_duplicate_removal = get('sampler_duplicate_removal')
        self._loader = overrides.get('loader', load_data)
        self._feat_idx_dic = overrides.get('feat_idx_dic')
        self.init()
        print('{} train graphs; {} validation graphs; {} test graphs'.format(
            self.train_data.num_graphs(), self.valid_data.num_graphs(), self.test_data.num_graphs()))

    def init(self):
        orig_train_data = self._loader(self.dataset, True)
        self.n = len(orig_train_data.graphs)
        train_gs, valid_gs = self._train_val_split(orig_train_data)
        test_gs = self._loader(self.dataset, False).graphs
        self.node_feat_encoder = self._get_node_feature_encoder(orig_train_data.graphs + test_gs)
        self._check_graphs_num(test_gs, 'test')
        mk = lambda gs: ModelGraphList(self.sampler, self.sample_num, self.sampler_duplicate_removal,  # noqa: E731
                                       gs, self.node_feat_encoder)
        self.train_data, self.valid_data, self.test_data = mk(train_gs), mk(valid_gs), mk(test_gs)
        self.m = self.test_data.num_graphs()
        assert len(train_gs) + len(valid_gs) == len(orig_train_data.graphs)

    def input_dim(self):
        return self.node_feat_encoder.input_dim()

    def get_graph_pair(self, train_val_test):
        return self._get_graph_collection(train_val_test).get_graph_pair()

    def get_graph_triple_for_hinge_loss(self, train_val_test):
        return self._get_graph_collection(train_val_test).get_triple_for_hinge_loss()

    def get_orig_train_graph(self, orig_train_id):
        trainlen, vallen = self.train_data.num_graphs(), self.valid_data.num_graphs()
        if 0 <= orig_train_id < trainlen:
            return self.train_data.get_graph(orig_train_id)
        if orig_train_id < trainlen + vallen:
            return self.valid_data.get_graph(orig_train_id - trainlen)
        assert False

    def get_dist(self, g1, g2, dist_calculator):
        return dist_calculator.calculate_dist(g1, g2)

    def m_n(self):
        return self.m, self.n

    def _train_val_split(self, orig_train_data):
        if self.valid_percentage < 0 or self.valid_percentage > 1:
            raise RuntimeError('valid_percentage {} must be in [0, 1]'.format(self.valid_percentage))
        gs = orig_train_data.graphs
        sp = int(len(gs) * (1 - self.valid_percentage))
        train_graphs, valid_graphs = gs[0:sp], gs[sp:]
        self._check_graphs_num(train_graphs, 'train')
        self._check_graphs_num(valid_graphs, 'validation')
        return train_graphs, valid_graphs

    def _check_graphs_num(self, graphs, label):
        if len(graphs) <= 2:
            raise RuntimeError('Insufficient {} graphs {}'.format(label, len(graphs)))

    def _get_node_feature_encoder(self, gs):
        if self.node_feat_encoder == 'onehot':
            return NodeFeatureOneHotEncoder(gs, self.node_feat_name, self._feat_idx_dic)
        raise RuntimeError('Unknown node_feat_encoder {}'.format(self.node_feat_encoder))

    def _get_graph_collection(self, train_val_test):
        if train_val_test == 'train':
            return self.train_data
        if train_val_test == 'val':
            return self.valid_data
        if train_val_test == 'test':
            return self.test_data
        raise RuntimeError('Unknown train_val_test {}'.format(train_val_test))
