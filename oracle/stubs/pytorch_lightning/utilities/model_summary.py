class ModelSummary(object):
    def __init__(self, model, *args, **kwargs):
        self.model = model

    def __str__(self):
        rows = ['  | Name            | Type            | Params']
        for i, (n, m) in enumerate(self.model.named_children()):
            rows.append(f'{i} | {n} | {type(m).__name__} | '
                        f'{sum(p.numel() for p in m.parameters())}')
        return '\n'.join(rows)
