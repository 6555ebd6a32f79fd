"""Process-global id allocation for boundary conditions (boundary_condition_registry.py:6-30)."""


class BoundaryConditionRegistry:
    def __init__(self):
        self.id_to_bc = {}
        self.bc_to_id = {}
        self.next_id = 1  # 0 = no boundary condition

    def register_boundary_condition(self, boundary_condition):
        new_id = self.next_id
        if new_id > 253:
            raise ValueError("bc_mask is a uint8 field: at most 253 boundary conditions per process (254/255 are reserved)")
        self.next_id += 1
        self.id_to_bc[new_id] = boundary_condition
        self.bc_to_id[boundary_condition] = new_id
        print(f"registered bc {boundary_condition} with id {new_id}")
        return new_id


boundary_condition_registry = BoundaryConditionRegistry()
