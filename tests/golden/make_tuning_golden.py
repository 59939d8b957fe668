"""Extracts the literal tables of the reference's sweep driver (model/Siamese/tuning.py:19-47: dataset, header,
model, the seven ranges) with ``ast`` -- the module imports TensorFlow, so it cannot be imported here -- and
writes them to tests/golden/tuning.json.  Run in the build container (needs /root/reference)."""
import ast
import json
import os

SRC = '/root/reference/model/Siamese/tuning.py'
WANT = ('dataset', 'header', 'model', 'dist_norm_range', 'yeta_range_norm', 'yeta_range_nonorm', 'final_act_range',
        'lr_range', 'iter_range', 'val_ratio_range', 'dropout_range')


def main():
    tree = ast.parse(open(SRC).read())
    out = {}
    for node in tree.body:
        if isinstance(node, ast.Assign) and len(node.targets) == 1 and isinstance(node.targets[0], ast.Name):
            name = node.targets[0].id
            if name in WANT:
                try:
                    out[name] = ast.literal_eval(node.value)
                except ValueError:      # iter_range = list(range(500, 2001, 500))
                    out[name] = eval(compile(ast.Expression(node.value), SRC, 'eval'), {'list': list, 'range': range})
    # the four layer strings tune() installs (tuning.py:74-93)
    layers = {}
    for node in ast.walk(tree):
        if isinstance(node, ast.Call) and getattr(node.func, 'id', '') == 'reset_flag' and len(node.args) == 4:
            key = node.args[2]
            if isinstance(key, ast.Constant) and str(key.value).startswith('layer_'):
                layers[key.value] = ast.literal_eval(node.args[3])
    out['layers'] = [layers[k] for k in sorted(layers)]
    assert set(WANT) <= set(out), sorted(set(WANT) - set(out))
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), 'tuning.json')
    json.dump(out, open(path, 'w'), indent=1, sort_keys=True)
    print('wrote', path)


if __name__ == '__main__':
    main()
