"""Enums of the hot path, numerically identical to the reference's (epidemiologyclasses.py:1-65, cellsclasses.py:1-19),
so that values written back into the reference's containers compare equal to the reference's own IntEnums."""
from enum import IntEnum


class SEIRState(IntEnum):
    Undefined = 0
    Susceptible = 1
    Exposed = 2
    Infectious = 3
    Recovered = 4
    Deceased = 5


class InfectionType(IntEnum):
    Undefined = 0
    PreAsymptomatic = 1
    PreSymptomatic = 2
    Asymptomatic = 3
    Symptomatic = 4


class Severity(IntEnum):
    Undefined = 0
    Mild = 1
    Severe = 2
    Critical = 3


class SEIRStateTransition(IntEnum):
    ExposedToInfectious = 0
    InfectiousToSymptomatic = 1
    SymptomaticToSevere = 2
    SevereToCritical = 3
    CriticalToDeath = 4
    AsymptomaticToRecovery = 5
    MildToRecovery = 6
    SevereToRecovery = 7
    CriticalToRecovery = 8
    RecoveredToSusceptible = 9


class QuarantineType(IntEnum):
    Symptomatic = 0
    Positive = 1
    PositiveContact = 2
    SecondaryContact = 3


class CellType(IntEnum):
    Household = 0
    Workplace = 1
    Accommodation = 2
    Hospital = 3
    Entertainment = 4
    School = 5
    Classroom = 6
    Institution = 7
    Transport = 8
    Religion = 9
    Airport = 10


class SimCellType(IntEnum):
    Residence = 0
    Workplace = 1
    School = 2
    Community = 3
