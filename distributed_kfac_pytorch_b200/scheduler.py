"""Epoch-stepped schedule for the K-FAC hyper-parameters.

Pure host logic kept with the reference's interface (kfac/scheduler.py:1-94 in
/root/reference): it mutates ``kfac.param_groups[0]`` in place."""


class KFACParamScheduler(object):
    """Multiplies ``damping`` by ``damping_alpha`` at every step listed in
    ``damping_schedule`` and the two update frequencies by ``update_freq_alpha`` at
    every step in ``update_freq_schedule``; call ``step()`` once per epoch."""

    def __init__(self, kfac, damping_alpha=1, damping_schedule=None, update_freq_alpha=1,
                 update_freq_schedule=None, start_step=0):
        self.kfac = kfac
        group = kfac.param_groups[0]
        self.damping_base = group['damping']
        self.damping_alpha = damping_alpha
        self.damping_schedule = damping_schedule
        self.damping_factor_func = self._make_factor_func(damping_schedule, damping_alpha)
        self.factor_update_freq_base = group['factor_update_freq']
        self.inv_update_freq_base = group['inv_update_freq']
        self.update_freq_alpha = update_freq_alpha
        self.update_freq_schedule = update_freq_schedule
        self.update_freq_factor_func = self._make_factor_func(update_freq_schedule,
                                                              update_freq_alpha)
        self._step = start_step

    @staticmethod
    def _make_factor_func(schedule, alpha):
        milestones = sorted(schedule, reverse=True) if schedule is not None else []
        if schedule is not None:
            schedule.sort(reverse=True)   # the reference sorts the caller's list too

        def factor(step):
            out = 1.
            for t in milestones:
                if step >= t:
                    out *= alpha
            return out
        return factor

    def state_dict(self):
        return {k: v for k, v in self.__dict__.items() if k != 'kfac' and 'func' not in k}

    def load_state_dict(self, state_dict):
        self.__dict__.update(state_dict)

    def step(self, step=None):
        self._step = step if step is not None else self._step + 1
        group = self.kfac.param_groups[0]
        group['damping'] = self.damping_base * self.damping_factor_func(self._step)
        f = self.update_freq_factor_func(self._step)
        group['factor_update_freq'] = int(self.factor_update_freq_base * f)
        group['inv_update_freq'] = int(self.inv_update_freq_base * f)
