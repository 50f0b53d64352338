"""Input domains of the Fairify experiments, as data.

Feature order is the order of the non-label DataFrame columns the reference loaders produce; every
hot-path function indexes the box *by column name in that order* (reference `utils/prune.py:108-114`,
`:206-216`).  The `range_dict`s are the module-level constants of the drivers; their key order is the
order `partition` / `partitioned_ranges` iterate in (`utils/input_partition.py:20,56`), which differs
from the column order (e.g. Bank's `age` is the first column but the last key).

Sources (reference file:line):
  AC  columns `utils/verif_utils.py:127-130` (minus fnlwgt), ranges `src/AC/Verify-AC.py:46-58`
  GC  columns `utils/verif_utils.py:195-201` (minus personal_status, plus `sex` appended by
      `utils/standard_data.py:50-53`), ranges `src/GC/Verify-GC.py:41-60`
  BM  columns `utils/verif_utils.py:311-313`, ranges `src/BM/Verify-BM.py:32-47`
  CP  columns `utils/verif_utils.py:246`, ranges `src/CP/Verify-CP.py:49-54`
"""
from __future__ import annotations

from collections import OrderedDict

AC_COLUMNS = ['age', 'workclass', 'education', 'education-num', 'marital-status', 'occupation',
              'relationship', 'race', 'sex', 'capital-gain', 'capital-loss', 'hours-per-week',
              'native-country']
AC_LABEL = 'income-per-year'
AC_RANGES = OrderedDict([
    ('age', [10, 100]), ('workclass', [0, 6]), ('education', [0, 15]), ('education-num', [1, 16]),
    ('marital-status', [0, 6]), ('occupation', [0, 13]), ('relationship', [0, 5]), ('race', [0, 4]),
    ('sex', [0, 1]), ('capital-gain', [0, 19]), ('capital-loss', [0, 19]), ('hours-per-week', [1, 100]),
    ('native-country', [0, 40]),
])

GC_COLUMNS = ['status', 'month', 'credit_history', 'purpose', 'credit_amount', 'savings', 'employment',
              'investment_as_income_percentage', 'other_debtors', 'residence_since', 'property', 'age',
              'installment_plans', 'housing', 'number_of_credits', 'skill_level', 'people_liable_for',
              'telephone', 'foreign_worker', 'sex']
GC_LABEL = 'credit'
GC_RANGES = OrderedDict([
    ('status', [0, 2]), ('month', [0, 80]), ('credit_history', [0, 2]), ('purpose', [0, 9]),
    ('credit_amount', [0, 20000]), ('savings', [0, 2]), ('employment', [0, 2]),
    ('investment_as_income_percentage', [1, 4]), ('other_debtors', [0, 2]), ('residence_since', [1, 4]),
    ('property', [0, 2]), ('age', [0, 1]), ('installment_plans', [0, 2]), ('housing', [0, 2]),
    ('number_of_credits', [1, 4]), ('skill_level', [0, 3]), ('people_liable_for', [1, 2]),
    ('telephone', [0, 1]), ('foreign_worker', [0, 1]), ('sex', [0, 1]),
])

BM_COLUMNS = ['age', 'job', 'marital', 'education', 'default', 'housing', 'loan', 'contact', 'month',
              'day_of_week', 'duration', 'emp.var.rate', 'campaign', 'pdays', 'previous', 'poutcome']
BM_LABEL = 'y'
BM_RANGES = OrderedDict([
    ('job', [0, 10]), ('marital', [0, 2]), ('education', [0, 6]), ('default', [0, 1]), ('housing', [0, 1]),
    ('loan', [0, 1]), ('contact', [0, 1]), ('month', [0, 11]), ('day_of_week', [0, 6]),
    ('emp.var.rate', [-3, 1]), ('duration', [0, 5000]), ('campaign', [1, 50]), ('pdays', [0, 999]),
    ('previous', [0, 7]), ('poutcome', [0, 2]), ('age', [0, 1]),
])

CP_COLUMNS = ['Two_yr_Recidivism', 'Number_of_Priors', 'Age', 'Race', 'Female', 'Misdemeanor']
CP_LABEL = 'score_factor'
CP_RANGES = OrderedDict([
    ('Two_yr_Recidivism', [0, 1]), ('Number_of_Priors', [0, 38]), ('Age', [0, 1]), ('Race', [0, 1]),
    ('Female', [0, 1]), ('Misdemeanor', [0, 1]),
])


class Domain:
    """One dataset's feature order, label column, full input range and default protected attribute."""

    def __init__(self, name, columns, label, ranges, pa, threshold):
        self.name = name
        self.columns = list(columns)
        self.label = label
        self.ranges = OrderedDict((k, list(v)) for k, v in ranges.items())
        self.pa = list(pa)
        self.threshold = threshold

    @property
    def d(self):
        return len(self.columns)

    def copy_ranges(self):
        return OrderedDict((k, list(v)) for k, v in self.ranges.items())


# default PA / PARTITION_THRESHOLD of the src/ drivers (`src/AC/Verify-AC.py:36,61`,
# `src/GC/Verify-GC.py:31,63`, `src/BM/Verify-BM.py:23,51`, `src/CP/Verify-CP.py:38,71`)
DOMAINS = {
    'AC': Domain('AC', AC_COLUMNS, AC_LABEL, AC_RANGES, ['sex'], 10),
    'GC': Domain('GC', GC_COLUMNS, GC_LABEL, GC_RANGES, ['age'], 100),
    'BM': Domain('BM', BM_COLUMNS, BM_LABEL, BM_RANGES, ['age'], 100),
    'CP': Domain('CP', CP_COLUMNS, CP_LABEL, CP_RANGES, ['Race'], 5),
}

# PARTITION_THRESHOLD of the scaled experiments (`stress/AC/Verify-AC.py:23`, `stress/GC/Verify-GC.py:31`,
# `stress/BM/Verify-BM.py:23`)
STRESS_THRESHOLD = {'AC': 6, 'GC': 10, 'BM': 10}


def domain_for_model(model_name: str) -> Domain:
    return DOMAINS[model_name.split('-')[0].lstrip('a')]
